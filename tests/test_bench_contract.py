"""bench.py's contract on a machine without a GPU: the reference arm (CPU restatement, bounded sample)
prints one JSON line with the agreed keys; the GPU arm refuses to run (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=600):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=timeout, cwd=ROOT)


def test_reference_arm_prints_the_contract_line(built_lib):
    # the full step is 2^26 queries (about a second on the GPU box's host cores); a CPU test takes 2^22
    r = _run("--impl", "reference", "--steps", "2", "--warmup", "1", "--log2-batch", "22")
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    # --steps / --warmup are honoured as given and a step is the whole batch (same config keys as the GPU arm)
    assert line["steps"] == 2 and line["warmup"] == 1 and line["config"]["queries_per_gpu_per_step"] == 1 << 22
    assert line["impl"] == "reference" and line["metric"] == "merl_eval_pdf_throughput" and line["unit"] == "Gquery/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"]


def test_gpu_arm_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    r = _run("--steps", "1", "--warmup", "1", timeout=300)
    assert r.returncode != 0 and "CUDA" in (r.stderr + r.stdout)


def test_reference_arm_ignores_torchrun_thread_cap(built_lib):
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm must still use every host thread
    of its affinity mask, so that the driver's ratio means the same at every N."""
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "1", "--log2-batch", "22"], capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0)) and line["n_gpus"] == 2
    env["RANK"] = "1"                          # the other ranks exit 0 without work
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_cpu_legs_subprocess_reports_config1(built_lib):
    sys.path.insert(0, ROOT)
    import bench
    c = bench.cpu_legs_subprocess()
    assert c["cores"] >= 1 and c["eval_pdf_all_threads"]["gquery_s"] > 0
    one, all_ = c["config1_2p20"]["single_thread"], c["config1_2p20"]["all_threads"]
    assert one["threads"] == 1 and all_["threads"] == c["cores"] and c["config1_2p20"]["pairs"] == 1 << 20
    for leg in (one, all_):
        assert leg["eval_pdf_gquery_s"] > 0 and leg["sample_gquery_s"] > 0
    # the render half of the metric (spp*Mpix/s) has its CPU figure too: configs 2 and 5 through the render oracle
    r = c["render"]
    assert r["cfg2_all_threads"]["spp_mpix_per_s"] > 0 and r["cfg2_all_threads"]["threads"] == c["cores"]
    assert r["cfg2_single_thread"]["threads"] == 1 and r["cfg5_all_threads"]["spp_mpix_per_s"] > 0
    assert (r["cfg5_all_threads"]["width"], r["cfg5_all_threads"]["height"]) == (1024, 1024)
