"""The header is plain C and the library is usable from a C program (no Python, no torch)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

C_PROGRAM = r"""
#include <stdio.h>
#include <string.h>
#include "dlb200.h"

int main(void) {
    char topo[4096];
    int count = -1;
    dlb_sgd_hyper_t h = {0.01f, 0.5f, 0.0f, 0.0f, 1.0f, 0};
    dlb_fused_sgd_t f;
    memset(&f, 0, sizeof f);
    f.hyper = h;
    if (dlb_version() < 100) return 1;
    if (dlb_emulated_flags_bytes(8) != 8u * (DLB_SIGNAL_BYTES + 4 * DLB_STATE_WORDS)) return 2;
    if (dlb_emulated_static_scratch_bytes() != dlb_emulated_flags_bytes(DLB_MAX_RANKS)) return 3;
    if (dlb_topology_json(topo, sizeof topo) != 0 || strstr(topo, "\"devices\"") == NULL) return 4;
    /* argument checking happens before any CUDA call, so it works without a GPU */
    if (dlb_pack(NULL, 99, NULL, NULL, NULL, NULL, 0, 1.0f, NULL) != DLB_E_ARG) return 5;
    if (strstr(dlb_last_error(), "dtype") == NULL) return 6;
    /* round-2 entry points: descriptor form of the allreduce, named pack kernels, host-side generator */
    {
        dlb_allreduce_desc_t d;
        dlb_emulated_desc_t e;
        float a[64], b[64];
        int64_t labels[16];
        int i, same = 1;
        memset(&d, 0, sizeof d);
        memset(&e, 0, sizeof e);
        if (dlb_allreduce(NULL, NULL) != DLB_E_ARG) return 7;
        d.algo = 17;
        if (dlb_allreduce(&d, NULL) != DLB_E_ARG || strstr(dlb_last_error(), "algo") == NULL) return 8;
        d.algo = DLB_ALGO_TWOSHOT_BULK;
        if (dlb_allreduce(&d, NULL) != DLB_E_ARG) return 9;          /* null bufs */
        if (dlb_allreduce_emulated_ex(&e, NULL) != DLB_E_ARG) return 10;
        if (dlb_pack_ex(NULL, DLB_F32, NULL, NULL, NULL, NULL, 0, 1.0f, 9, NULL) != DLB_E_ARG) return 11;
        if (dlb_host_fill_uniform(a, DLB_F32, 64, 5, 1, 1) != 0 || dlb_host_fill_uniform(b, DLB_F32, 64, 5, 1, 3) != 0) return 12;
        for (i = 0; i < 64; ++i) same = same && a[i] == b[i] && a[i] >= 0.0f && a[i] < 1.0f;
        if (!same) return 13;                                        /* independent of the thread count */
        if (dlb_host_fill_uniform(b, DLB_F32, 64, 5, 2, 1) != 0 || memcmp(a, b, sizeof a) == 0) return 14;
        if (dlb_host_fill_randint(labels, 16, 10, 5, 1) != 0) return 15;
        for (i = 0; i < 16; ++i) if (labels[i] < 0 || labels[i] >= 10) return 16;
        if (dlb_host_fill_uniform(a, DLB_F16, 64, 5, 1, 1) != DLB_E_ARG) return 17;
    }
    (void)dlb_device_count(&count);
    printf("ok version=%d devices=%d sgd_args=%zu bytes\n", dlb_version(), count, sizeof f);
    return 0;
}
"""


@pytest.mark.skipif(shutil.which("gcc") is None, reason="no C compiler")
def test_header_is_c_and_library_links_from_c(tmp_path):
    from distributed_learning_b200.ops import _lib

    _lib.lib()                                   # builds the library if it is missing
    libdir = os.path.dirname(_lib.library_path())
    src = tmp_path / "abi.c"
    src.write_text(C_PROGRAM)
    exe = tmp_path / "abi"
    cc = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                         "-L", libdir, "-ldlb200", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True)
    assert run.returncode == 0, (run.returncode, run.stdout, run.stderr)
    assert run.stdout.startswith("ok version=")
