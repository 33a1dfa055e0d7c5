"""``python -m distributed_learning_b200 n r d D a i M R g [options]`` -- same as ``main.py``."""
import sys

from .launcher import main

if __name__ == "__main__":
    sys.exit(main())
