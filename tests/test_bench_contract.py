"""bench.py's host-side contract, checked without a GPU: the CPU arm (`--impl reference`) prints one JSON line with
the keys the driver parses, uses every host thread even when launched with OMP_NUM_THREADS=1 (what
torch.distributed.run exports), reports on the GPU arm's own `config`, and ranks other than 0 exit without work."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env_extra):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(REPO, "bench.py")] + args, cwd=REPO, env=env,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)


def test_reference_arm_line_and_threads():
    r = _run(["--impl", "reference", "--workload", "mnist4096", "--gpus", "2", "--steps", "2", "--warmup", "1",
              "--no-extra"], {"OMP_NUM_THREADS": "1", "RANK": "0", "WORLD_SIZE": "2"})
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["unit"] == "samples/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    threads = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == threads   # not the 1 thread torchrun's environment asks for
    # the same config object the GPU arm prints for this workload and GPU count
    sys.path.insert(0, REPO)
    import bench
    assert line["config"] == bench.line_config("mnist4096", 2)
    assert line["config"]["workload"] == "mnist4096" and line["vs_baseline"] is None


def test_other_ranks_of_the_reference_arm_do_nothing():
    r = _run(["--impl", "reference", "--gpus", "2"], {"RANK": "1", "WORLD_SIZE": "2"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_workload_tables_are_consistent():
    sys.path.insert(0, REPO)
    import bench
    assert bench.DEFAULT_WORKLOAD == "wide8x16384"
    for name, (dims, acts, rows, dtype, micro) in bench.WORKLOADS.items():
        assert len(acts) == len(dims) - 1 and rows % micro == 0 and dtype in ("f32", "bf16")
        cfg = bench.line_config(name, 1)
        assert cfg["rows_per_step"] == rows and cfg["micro_batches"] == micro
    # SURVEY.md section 8d figures
    assert bench.flops_per_sample([784, 128, 64, 10]) == 454400
    assert bench.flops_per_sample([8192] * 5) == 11 * 2 * 8192 ** 2
    assert bench.flops_per_sample([16384] * 9) == 23 * 2 * 16384 ** 2
    assert bench.bytes_per_sample([784, 256, 256, 10], 4) == 18684
    assert bench.PIPE_DW_CHUNK_ROWS % (bench.WORKLOADS["wide8x16384"][2] // bench.WORKLOADS["wide8x16384"][4]) == 0
