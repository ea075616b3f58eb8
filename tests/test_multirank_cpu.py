"""The N > 1 host logic on CPU (world_size 2, gloo): shard bounds + global-index weights + one all-gather of the 336-byte
records reproduce the single-call pairing operands; and the reference arm of bench.py under torchrun prints exactly one
JSON line from rank 0 while the other rank exits 0 without work."""
import json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(port, *args):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", str(port)]
    return subprocess.run(cmd + list(args), capture_output=True, text=True, timeout=600, cwd=ROOT)


def test_shard_bounds_cover_the_batch():
    from kateth_b200.setup import shard_bounds
    for n in (0, 1, 5, 4096, 16384, 65536 + 3):
        for world in (1, 2, 4, 8):
            b = [shard_bounds(n, world, g) for g in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[g][1] == b[g + 1][0] for g in range(world - 1))


def test_sharded_batch_verification_over_gloo():
    p = _torchrun(29631, os.path.join(ROOT, "tests", "_gloo_worker.py"))
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "rank 0 of 2" in p.stdout and "rank 1 of 2" in p.stdout


def test_reference_arm_under_torchrun_prints_one_line():
    p = _torchrun(29632, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--ref-blobs", "2")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["cpu_baseline"]["kind"] == "port" and j["e2e"]["h2d_bytes_per_step"] == 0 and j["value"] > 0
