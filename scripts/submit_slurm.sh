#!/bin/bash
# SLURM template (replaces the reference's submit.sh / submit_cpu.sh / submits/*.sh family).
# One task per GPU on ONE node; rank and size reach main.py through envarg://, exactly as the
# reference's mpirun launches did (OMPI_COMM_WORLD_* / PMI_*).  Edit the four variables.
#SBATCH -J dlb200
#SBATCH --nodes=1
#SBATCH --ntasks-per-node=8
#SBATCH --gres=gpu:8
#SBATCH --cpus-per-task=12
#SBATCH --time=00:30:00
#SBATCH --no-requeue
GPUS=${SLURM_NTASKS_PER_NODE:-8}
MODEL=${MODEL:-imagenet}
EXPERIMENT=${EXPERIMENT:-experiment2}
BACKEND=${BACKEND:-b200}
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
options="envarg://SLURM_NTASKS envarg://SLURM_PROCID 1 $GPUS 127.0.0.1 lo $MODEL ${DATA_ROOT:-/tmp/dlb200_data} 1 \
 --job_id ${SLURM_JOB_ID:-job} --experiment $EXPERIMENT --backend $BACKEND --limit_batches 30 --random_input 1"
echo "JobID: ${SLURM_JOB_ID:-none}  nodes: ${SLURM_JOB_NODELIST:-local}  command: python main.py $options"
# srun exports SLURM_PROCID/SLURM_NTASKS per task; with OpenMPI use
#   mpirun -npernode $GPUS python main.py envarg://OMPI_COMM_WORLD_SIZE envarg://OMPI_COMM_WORLD_RANK ...
srun python "$ROOT/main.py" $options
