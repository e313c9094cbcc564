"""The drop-in executables (host/*.out, C on top of libbbgpu.so): command line, validation and output formats of the
reference's pfsp_multigpu_cuda.c / nqueens_multigpu_cuda.cu. CPU part: flag validation
and the settings block; GPU part: whole runs whose printed counts must equal the reference's."""
import os
import re
import subprocess

import pytest

from __graft_entry__ import PKG_DIR

PFSP = os.path.join(PKG_DIR, "host", "pfsp_multigpu_cuda.out")
NQ = os.path.join(PKG_DIR, "host", "nqueens_multigpu_cuda.out")


def run(exe, *args, cwd=None):
    return subprocess.run([exe, *map(str, args)], capture_output=True, text=True, cwd=cwd, timeout=600)


def test_executables_are_built():
    assert os.path.exists(PFSP) and os.path.exists(NQ), "run __graft_entry__.build()"


@pytest.mark.parametrize("args,msg", [
    (["-i", 0], "Error: unsupported Taillard's instance"), (["-i", 121], "Error: unsupported Taillard's instance"),
    (["-l", 3], "Error: unsupported lower bound function"), (["-u", 2], "Error: unsupported upper bound initialization"),
    (["-m", 0], "Error: unsupported minimal pool for GPU initialization"),
    (["-m", 30, "-M", 20], "Error: unsupported maximal pool for GPU initialization"),
    (["-T", 3], "Error: unsupported maximal pool for CPU multi-core"),
    (["-w", 2], "Error: unsupported Intra-node Work Stealing option"),
    (["-p", 0], "Error: unsupported WS percentage for popFrontBulkFree"),
])
def test_pfsp_flag_validation_matches_reference(args, msg, tmp_path):
    r = run(PFSP, *args, cwd=tmp_path)  # PFSP_lib.c:216-313: message on stderr, EXIT_FAILURE, nothing else
    assert r.returncode == 1 and r.stderr.strip() == msg and r.stdout == ""


def test_pfsp_settings_block_and_refusals(tmp_path):
    r = run(PFSP, "-C", 2, cwd=tmp_path)  # pfsp_multigpu_cuda.c:523-527
    assert r.returncode == 1 and r.stdout.startswith("C is set to 2. Invalid option for this version.")
    r = run(PFSP, "--inst", 111, "--lb", 0, "--ub", 0, "-C", 0, "-w", 0, cwd=tmp_path)  # 500 jobs: refused loudly
    assert r.returncode == 1 and "supports at most 200 x 20" in r.stderr
    assert r.stdout == ("\n=================================================\n"
                        "Multi-core Multi-GPU C+OpenMP+CUDA (1 GPU(s) - [0] Multi-core - [0] Work Stealing)\n\n"
                        "Resolution of PFSP Taillard's instance: ta111 (m = 20, n = 500)\n"
                        "Initial upper bound: inf\nLower bound function: lb1_d\nBranching rule: fwd\n"
                        "=================================================\n")


def test_nqueens_flag_validation(tmp_path):
    for args in (["-N", 0], ["-g", -1], ["-D", 0]):
        r = run(NQ, *args, cwd=tmp_path)  # nqueens_multigpu_cuda.cu:38-41
        assert r.returncode == 1 and r.stdout == "All parameters must be positive integers.\n"
    r = run(NQ, "-m", 30, "-M", 20, cwd=tmp_path)
    assert r.returncode == 1 and "greater or equal to m" in r.stderr
    r = run(NQ, "-N", 21, cwd=tmp_path)
    assert r.returncode == 1 and "must not exceed 20" in r.stderr


def parse_final(out):
    tail = out
    tree = int(re.findall(r"Size of the explored tree: (\d+)", tail)[-1])
    sol = int(re.findall(r"Number of explored solutions: (\d+)", tail)[-1])
    mk = re.findall(r"Optimal makespan: (\d+)", tail)
    return tree, sol, int(mk[-1]) if mk else None


@pytest.mark.gpu
@pytest.mark.parametrize("args,exp", [(["-i", 14, "-l", 1], (2573652, 2648, 1377)), (["-i", 14, "-l", 0], (2573652, 2648, 1377)),
                                      (["-i", 14, "-l", 2, "-M", 100000], (144639, 0, 1377)),
                                      (["-i", 2, "-l", 2, "-C", 0], (7, 0, 1359)),
                                      # 50-, 100- and 200-job classes: what the reference rebuilt with MAX_JOBS 50 / 100 / 200 prints
                                      (["-i", 32, "-l", 2, "-C", 0], (3, 0, 2834)),
                                      (["-i", 66, "-l", 0, "-C", 0], (11, 0, 5135)),
                                      (["-i", 94, "-l", 2, "-C", 0], (55, 0, 10889))])
def test_pfsp_run_prints_reference_counts_and_csv(args, exp, tmp_path, golden):
    r = run(PFSP, *args, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert parse_final(r.stdout) == exp
    for block in ("Initial search on CPU completed", "Search on Parallel GPU completed", "Final on CPU completed",
                  "Exploration terminated."):
        assert block in r.stdout
    assert re.search(r"Elapsed time: \d+\.\d{4} \[s\]\n=+\n$", r.stdout)
    rows = open(os.path.join(tmp_path, "multigpu.csv")).read().splitlines()
    assert rows[0].startswith("instance_id,D,C,lower_bound,work_stealing,optimum,m,M,T,total_time,total_tree,total_sol,")
    assert rows[0].count(",") == 24 and len(rows) == 2
    f = rows[1].split(",")
    assert int(f[0]) == int(args[1]) and int(f[10]) == exp[0] and int(f[11]) == exp[1] and int(f[5]) == exp[2]
    assert rows[1].count('"[') == 13


@pytest.mark.gpu
def test_nqueens_run_prints_reference_counts(tmp_path, golden):
    r = run(NQ, "-N", 12, "-g", 2, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert parse_final(r.stdout)[:2] == (856188, 14200)
    assert "Resolution of the 12-Queens instance\n  with 2 safety check(s) per evaluation\n" in r.stdout
    assert "Search on GPU completed" in r.stdout and "Exploration terminated." in r.stdout


# ------------------------------------------------------------------ bench.py contract

def _bench(*args):
    import json
    import sys
    from __graft_entry__ import ROOT
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *map(str, args)], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout  # exactly one JSON line on stdout
    return json.loads(lines[0])


def test_bench_reference_arm_runs_without_a_gpu():
    """`bench.py --impl reference` times the reference's own CPU code (oracle/_ref) on a bounded sample"""
    d = _bench("--impl", "reference", "--workload", "pfsp:14:2", "--steps", 1, "--warmup", 0, "--ref-seconds", 1)
    assert d["impl"] == "reference" and d["metric"] == "explored_nodes_per_sec" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


@pytest.mark.gpu
def test_bench_line_has_the_contract_keys():
    d = _bench("--workload", "pfsp:14:2", "--steps", 2, "--warmup", 3, "--ref-seconds", 1)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["config"]["tree_nodes"] == 144639 and d["gpu_launches"] > 0 and d["value"] > 0
    assert d["roofline"]["bound"] == "hbm" and 0 < d["roofline"]["frac"] < 1.5
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
