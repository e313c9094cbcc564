#!/usr/bin/env python
"""Turn the stdout of the reference's own binaries (kept by scripts/gpu_pin_reference.sh under gpurun_out/pins/ and
by the CPU runs under gpurun_out/cpu_pins/) into tests/golden/ref_gpu_runs.json.

Every entry is a run of UNMODIFIED reference code: `ref_pfsp_multigpu_cuda.out` / `ref_nqueens_multigpu_cuda.out` are
pfsp/pfsp_multigpu_cuda.c and nqueens/nqueens_multigpu_cuda.cu rebuilt for sm_100a (make -C oracle refgpu) and run on
a B200; `nqueens_c.out` / `pfsp_omp_c.out` are the reference's CPU programs (make -C oracle ref) run in the build
container. The numbers parsed are the ones the reference prints at pfsp_multigpu_cuda.c:503-510 /
nqueens_multigpu_cuda.cu:364-369 / nqueens_c.c:150-155. bench.py and the -m gpu tests read this file.

    python scripts/pins_to_golden.py            # merges whatever outputs are present into the JSON
"""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "ref_gpu_runs.json")


def parse(path):
    txt = open(path).read()
    cmd = re.search(r"^\$ (.*)$", txt, re.M)
    trees = re.findall(r"Size of the explored tree: (\d+)", txt)
    sols = re.findall(r"Number of explored solutions: (\d+)", txt)
    mk = re.findall(r"Optimal makespan: (\d+)", txt)
    el = re.findall(r"Elapsed time: ([0-9.]+) \[s\]", txt)
    rc = re.search(r"^rc=(\d+)", txt, re.M)
    if not trees or (rc and rc.group(1) != "0"):
        return None
    r = {"command": cmd.group(1) if cmd else os.path.basename(path), "tree": int(trees[-1]), "sol": int(sols[-1]),
         "elapsed_s": float(el[-1]), "phase_elapsed_s": [float(x) for x in el[:-1]]}
    if mk:
        r["makespan"] = int(mk[-1])
    return r


def main():
    data = json.load(open(OUT)) if os.path.exists(OUT) else {"runs": {}}
    data["what"] = ("stdout of the reference's own programs, unmodified sources; GPU programs rebuilt for sm_100a and "
                    "run on the B200 box by scripts/gpu_pin_reference.sh, CPU programs run in the build container; "
                    "parsed by scripts/pins_to_golden.py")
    smi = os.path.join(ROOT, "gpurun_out", "pins", "nvidia_smi_L.txt")
    if os.path.exists(smi):
        data["nvidia_smi_L"] = open(smi).read().strip().splitlines()
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "pins", "ref_*.txt"))):
        r = parse(path)
        if not r:
            continue
        m = re.search(r"-i (\d+) -l (\d+)", r["command"])
        q = re.search(r"-N (\d+)", r["command"])
        d = re.search(r"-D (\d+)", r["command"])
        r["where"] = "B200, sm_100a rebuild of the reference's GPU driver, D=%s" % (d.group(1) if d else "1")
        key = ("pfsp:%s:%s" % (m.group(1), m.group(2))) if m else ("nq:%s" % q.group(1))
        data["runs"][key] = r
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "cpu_pins", "*.log"))):
        txt = open(path).read()
        name = os.path.basename(path)
        r = parse(path)
        if not r:
            continue
        if name.startswith("nqueens_c_N"):
            N = int(re.search(r"N(\d+)", name).group(1))
            r["command"] = "nqueens_c.out -N %d -g 1" % N
            r["where"] = "build container CPU, 1 thread (nqueens/nqueens_c.c, unmodified)"
            data["runs"]["nq:%d" % N] = r
        elif name.startswith("pfsp_omp_c_ta"):
            m = re.search(r"ta(\d+)_lb(1d|1|2)", name)
            lb = {"1d": 0, "1": 1, "2": 2}[m.group(2)]
            r["command"] = "pfsp_omp_c.out -i %d -l %d -u 1 -C 5" % (int(m.group(1)), lb)
            r["where"] = "build container CPU, 5 threads (pfsp/pfsp_omp_c.c, unmodified)"
            data["runs"]["pfsp:%d:%d" % (int(m.group(1)), lb)] = r
        r["started_finished_utc"] = re.findall(r"^\d{4}-\d\d-\d\dT[\d:]+Z$", txt, re.M)
    json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)
    for k, v in sorted(data["runs"].items()):
        print(k, v["tree"], v["sol"], v.get("makespan"), v["elapsed_s"], "|", v["command"])


if __name__ == "__main__":
    main()
