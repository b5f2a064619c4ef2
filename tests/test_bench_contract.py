"""CPU: the bench.py contract the driver relies on for the reference arm — ONE JSON line on stdout with the agreed
keys, rank 0 only under a multi-rank launch.  (The GPU arm needs a B200 and is exercised by the driver.)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "vit_s16",
                           "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env,
                          cwd=ROOT)


def test_reference_arm_prints_one_json_line_with_contract_keys():
    # torch.distributed.run exports OMP_NUM_THREADS=1 to its children: the arm must reset the thread count itself
    r = _run({"RANK": "0", "WORLD_SIZE": "1", "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["unit"] == "images/s" and j["higher_is_better"] is True
    for key in ("metric", "value", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "vs_baseline", "dtype", "data",
                "config", "cpu_baseline", "e2e"):
        assert key in j, key
    assert j["value"] > 0 and j["e2e"]["value"] == j["value"]
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["e2e"]["d2h_bytes_per_step"] == 0
    from oracle import ref_loader
    assert j["cpu_baseline"]["kind"] == ("reference" if ref_loader.available() else "port")
    ncpu = min(os.cpu_count(), len(os.sched_getaffinity(0)))
    assert j["cpu_baseline"]["cores"] == ncpu, (j["cpu_baseline"], ncpu)   # every host core, not torchrun's 1
    assert "workload" in j["config"] and "model" not in j["config"]


def test_reference_arm_other_ranks_exit_silently():
    r = _run({"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""
