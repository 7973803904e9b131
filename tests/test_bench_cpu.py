"""bench.py's CPU side: the reference arm runs without the product library, reports one-thread and all-cores numbers for
both passes of benchmark.f90, and the host query generator is the one the device generator mirrors."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    env = dict(os.environ, OMP_NUM_THREADS="1")          # what torchrun exports: the arm must ignore it
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", *args], capture_output=True, text=True,
                         env=env, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


def test_reference_arm_c1_benchmark_f90():
    d = _run("--workload", "C1", "--steps", "1", "--warmup", "0")
    cb = d["cpu_baseline"]
    assert d["impl"] == "reference" and d["unit"] == "queries/s" and d["higher_is_better"] is True
    assert cb["kind"] == "port" and cb["cores"] == len(os.sched_getaffinity(0)) and cb["serial"]["cores"] == 1
    assert cb["value"] > 0 and cb["serial"]["value"] > 0 and cb["second_pass_with_hints"]["value"] > 0
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    assert "triangle" in d["config"]["workload"]


def test_reference_arm_small_c3_c4_c5_and_no_product_library():
    for wl in ("C3", "C4", "C5"):
        d = _run("--workload", wl, "--mesh-n", "10", "--steps", "1", "--warmup", "0", "--cpu-seconds", "0.5")
        assert d["value"] > 0 and d["cpu_baseline"]["serial"]["value"] > 0, wl
    # the arm must not map libiu_b200.so: run it under a loader trace
    code = ("import sys; sys.argv=['bench.py','--impl','reference','--workload','C3','--mesh-n','8','--steps','1','--warmup','0','--cpu-seconds','0.3'];"
            "import runpy; runpy.run_path(%r, run_name='__main__');"
            "maps=open('/proc/self/maps').read(); assert 'libiu_oracle.so' in maps; assert 'libiu_b200.so' not in maps, 'product library mapped'" % os.path.join(ROOT, "bench.py"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]


def test_config_is_the_same_in_both_arms():
    sys.path.insert(0, ROOT)
    import bench
    wl = bench.Workload("C3", 6).mesh()
    assert bench.config_of(wl) == bench.config_of(bench.Workload("C3", 6).mesh())
    per_q = wl.algorithmic_bytes(10)[1]
    assert per_q == 24 + 4 + 32                       # r + i_cell + 4 values
    assert bench.Workload("C5", 6).mesh().algorithmic_bytes(1)[1] == 24 + 4 + 64 + 8
    r = wl.host_queries(1000)
    assert r.shape == (1000, 3) and (r >= wl.qlo).all() and (r <= wl.qhi).all()
    assert np.array_equal(wl.host_queries(10, start=5), r[5:15])
