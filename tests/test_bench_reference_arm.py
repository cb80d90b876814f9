"""bench.py --impl reference: the CPU arm the driver times beside ours (host code only, no GPU).

The reference's CPU kernels for the path are single-threaded loops; at N > 1 the arm runs the reference's N-replica
data-parallel job, one host thread per replica (gather of every replica's SelectedRows gradient, applied by every
replica — framework/details/reduce_op_handle.cc:88-94, reduce_and_gather.h:56-93)."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_replica_job_applies_the_gathered_gradient_on_every_replica():
    import bench
    res = bench.time_cpu_job((500, 8, 64, 1.1, "sgd"), replicas=3, steps=2, warmup=1)
    assert res["cores"] == 3 and res["kind"] in ("reference", "port")
    assert res["value"] > 0 and res["ms_per_step"] > 0
    assert "3-replica job" in res["sample"] and "192 rows" in res["sample"]


def test_table_in_place_matches_the_value_range():
    import bench
    t = bench._table_in_place(1000, 4, seed=3)
    assert t.dtype == np.float32 and t.shape == (1000, 4)
    assert float(t.min()) >= -0.05 and float(t.max()) <= 0.05 and float(t.std()) > 0.02


def test_reference_arm_line_single_process():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 1 and line["unit"] == "ids/s"
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["cores"] == 1 and line["cpu_baseline"]["value"] == line["value"]
