"""The drop-in boundary end to end: the reference's UNMODIFIED pfsp_multigpu_cuda.c (host loop, pool,
generate_children, statistics - compiled where it lies by `make -C oracle refdropin`) linked against libbbgpu.so
instead of its own PFSP_gpu_lib.o, so that evaluate_gpu / lb1_alloc_gpu / lb2_alloc_gpu (pfsp_multigpu_cuda.c:190,
194, 295) resolve to this repo's compat kernels. Its printed counts must be the reference's. The binary only exists
where /root/reference was available at build time; it travels to the GPU box inside oracle/_ref/."""
import os
import re
import subprocess

import pytest

import orc

pytestmark = pytest.mark.gpu
EXE = os.path.join(orc.REF_DIR, "ref_pfsp_multigpu_with_bbgpu.out")


def parse(out):
    return (int(re.findall(r"Size of the explored tree: (\d+)", out)[-1]),
            int(re.findall(r"Number of explored solutions: (\d+)", out)[-1]),
            int(re.findall(r"Optimal makespan: (\d+)", out)[-1]))


@pytest.mark.parametrize("args,exp", [(["-i", "14", "-l", "1"], (2573652, 2648, 1377)),
                                      (["-i", "14", "-l", "0"], (2573652, 2648, 1377)),
                                      (["-i", "14", "-l", "2"], (144639, 0, 1377)),
                                      (["-i", "29", "-l", "2", "-M", "100000"], (9499307, 0, 2237))])
def test_unmodified_reference_driver_runs_on_our_kernels(args, exp, tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/ref_pfsp_multigpu_with_bbgpu.out not built (needs /root/reference at build time)")
    r = subprocess.run([EXE, *args, "-D", "1", "-C", "0"], capture_output=True, text=True, cwd=tmp_path, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert parse(r.stdout) == exp
