#!/usr/bin/env python3
"""bench.py -- blobs/s of kateth's EIP-4844 hot path on B200 (commit + blob-proof + batch-verify), see DESIGN.md "Measurement".

  python bench.py --gpus N --steps K --warmup W            the sm_100a engine (libkateth_b200.so, through its C ABI)
  python bench.py --impl reference --gpus N ...            kateth's CPU path as restated by oracle/ (kateth itself is Rust
                                                           over blst: neither toolchain nor crate exists here), all host threads

One "step" = one pass of the hot path over one batch of synthetic canonical-field-element blobs:
    blob_to_commitment (4096-point MSM)  ->  compute_blob_kzg_proof (challenge, barycentric evaluation, quotient, 4096-point MSM)
    ->  verify_blob_kzg_proof_batch (challenge, evaluation, G1 decompression, RLC sums, [all-gather of the 336-byte partials],
    host two-pairing check).
`value` times the step with inputs resident in HBM; `e2e` times the same step through the host-buffer C-ABI calls with the
blobs in pinned host memory (H2D of the blobs and D2H of commitments/proofs/verdict inside the timed region).
Under torchrun every rank owns `--blobs` blobs (weak scaling, the default) or `--global-blobs` / N of them (`--scaling strong`:
BASELINE configs[3]'s 16,384-blob batch cut over the GPUs).  The multi-GPU path is the C ABI's own: every rank joins the library's
group (kb200_group_create_rank; the unique id travels over torch.distributed), and batch verification is ONE
kb200_group_verify_blob_proof_batch* call per rank -- shard partial, ncclAllGather of the 336-byte records, host pairing, all inside
the library.  Prints ONE JSON line on rank 0.
"""
import argparse, json, os, subprocess, sys, threading, time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
SETUP_JSON = os.path.join(ROOT, "tests", "golden", "trusted_setup_4096.json")
BLOB = 131072
R_MOD = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
# SURVEY.md 8(d) fixed yardstick: one 4096-point, 255-bit MSM = 2.80e8 MAC32 (32x32->64 multiply-accumulates);
# an Fp product = 300 MAC32.  The table MSM gathers 4096 * windows affine points of 96 bytes.
MAC32_PER_MSM = 2.80e8
METRIC = "blobs/sec commit+prove+batch-verify"
METRIC_PAZ = "blobs/sec compute_kzg_proof+verify_kzg_proof (random points)"


def synth_blobs(n, seed):
    """n blobs of canonical field elements, uniform in [0, r): 256 random bits, top bit cleared, rejected if >= r
    (mirrors the distribution of Blob::random, /root/reference/src/blob.rs:66-76).  Returns uint8 [n*131072]."""
    import numpy as np
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(n * 4096, 32), dtype=np.uint8)
    a[:, 0] &= 0x7f
    rb = np.frombuffer(R_MOD.to_bytes(32, "big"), dtype=np.uint8)
    while True:
        # lexicographic a >= r on big-endian bytes: top byte > 0x73 fails, == 0x73 needs the full comparison
        bad = list(np.nonzero(a[:, 0] > rb[0])[0])
        bad += [i for i in np.nonzero(a[:, 0] == rb[0])[0] if bytes(a[i]) >= bytes(rb)]
        if not bad:
            break
        fresh = rng.integers(0, 256, size=(len(bad), 32), dtype=np.uint8)
        fresh[:, 0] &= 0x7f
        a[bad] = fresh
    return a.reshape(-1)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], None, [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2]); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "power_w_max": max(pw) if pw else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def env_ranks():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------ CPU legs (oracle/)
def cpu_leg(blobs_u8, n, threads):
    """kateth's CPU path (oracle restatement: Pippenger MSMs, per-term inversions, naive lincombs in batch verify) on n blobs
    with `threads` host threads, sharded over blobs.  Returns seconds for commit+prove and for batch-verify."""
    from oracle import oracle as O
    s = cpu_leg.setup = getattr(cpu_leg, "setup", None) or O.Setup.load_json(SETUP_JSON)
    cat = blobs_u8 if isinstance(blobs_u8, bytes) else bytes(blobs_u8[: n * BLOB])
    t0 = time.perf_counter()
    comm, proofs = s.commit_prove_many(cat, n, 2, threads)
    t1 = time.perf_counter()
    ok = s.verify_many(cat, comm, proofs, n, threads)
    t2 = time.perf_counter()
    assert ok, "oracle batch verification of its own proofs failed"
    return t1 - t0, t2 - t1, comm, proofs


def run_reference(args):
    rank, world, _ = env_ranks()
    if rank != 0:
        return
    threads = host_threads()
    n = max(threads, 8) if args.ref_blobs <= 0 else args.ref_blobs        # one blob per thread per step: a few seconds per step
    blobs = synth_blobs(n, 0x4b5a4700 + 3)
    times = []
    for it in range(args.warmup + args.steps):
        a, b, _, _ = cpu_leg(blobs, n, threads)
        if it >= args.warmup:
            times.append(a + b)
    dt = sum(times) / len(times)
    val = n / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "blobs/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 limbs (Fr 4x64, Fp 6x64 Montgomery)",
            "data": "synthetic canonical field elements, uniform in [0, r)",
            "config": {"workload": "commit + compute_blob_kzg_proof + verify_blob_kzg_proof_batch, 4096-element blobs (BASELINE configs[2]+[3] shape)",
                       "blobs_per_step": n, "note": "kateth is Rust over blst (no toolchain, crate absent): this is oracle/'s C restatement of its algorithm"},
            "cpu_baseline": {"value": val, "unit": "blobs/s", "cores": threads, "kind": "port", "sample": "%d blobs per step, %d threads" % (n, threads)},
            "e2e": {"value": val, "unit": "blobs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------ the sm_100a engine
def run_engine(args):
    import torch
    import torch.distributed as dist
    import kateth_b200 as K
    rank, world, local = env_ranks()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.blobs
    if args.scaling == "strong":
        if args.global_blobs % world:
            raise SystemExit("bench.py: --global-blobs must be a multiple of the number of GPUs")
        n = args.global_blobs // world
    s = K.Setup.load_json(SETUP_JSON, device=local, window_bits=args.window_bits)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    # the library's own multi-GPU group: rank 0's NCCL unique id goes to the other ranks over the torch process group
    uid = [K.Group.unique_id() if (rank == 0 and world > 1) else None]
    if world > 1:
        dist.broadcast_object_list(uid, src=0)
    g = K.Group.join(s, rank, world, uid[0])
    info0 = s.info()
    peak_wide, peak_fp = K.test_hooks.mac_peak(16384) if rank == 0 else (0.0, 0.0)

    h_blobs = torch.from_numpy(synth_blobs(n, 0x4b5a4700 + 3 + 1000 * rank)).pin_memory()
    h_comm = torch.empty(n * 48, dtype=torch.uint8).pin_memory()
    h_proof = torch.empty(n * 48, dtype=torch.uint8).pin_memory()
    d_blobs = h_blobs.to(dev)
    d_comm = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_proof = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(n, dtype=torch.int32, device=dev)

    def check(rc, what):
        if rc != 0:
            raise SystemExit("bench.py: %s returned status %d (%s)" % (what, rc, K.lib().kb200_last_error().decode()))

    stage_ev = []
    paz = args.workload == "proof_at_z"
    if paz:
        # evaluation points: canonical scalars from the blob generator's stream 1; commitments are computed once, untimed
        h_z = torch.from_numpy(synth_blobs((n + 4095) // 4096, 0x4b5a4700 + 5 + 1000 * rank)[: n * 32].copy()).pin_memory()
        h_y = torch.empty(n * 32, dtype=torch.uint8).pin_memory()
        d_z = h_z.to(dev)
        d_y = torch.empty(n * 32, dtype=torch.uint8, device=dev)
        check(s.blob_to_commitment_device(d_blobs.data_ptr(), n, d_comm.data_ptr(), d_status.data_ptr()), "blob_to_commitment")
        h_comm.copy_(d_comm)

    def verdict(rc, ok, what, want=True):
        check(rc, what)
        if ok is not want:
            raise SystemExit("bench.py: %s returned %s, expected %s" % (what, ok, want))

    def step_resident_paz(record):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if record else None
        if record: ev[0].record()
        check(s.proof_device(d_blobs.data_ptr(), d_z.data_ptr(), n, d_proof.data_ptr(), d_y.data_ptr(), d_status.data_ptr()), "compute_proof")
        if record: ev[1].record()
        verdict(*g.verify_points_device([d_proof.data_ptr()], [d_comm.data_ptr()], [d_z.data_ptr()], [d_y.data_ptr()], [n], rank * n, world * n, [d_status.data_ptr()]),
                "group_verify_proof_batch_device")
        if record: ev[2].record(); stage_ev.append(ev)

    def step_e2e_paz():
        check(s.proof_host(h_blobs.data_ptr(), h_z.data_ptr(), n, h_proof.data_ptr(), h_y.data_ptr()), "compute_proof(host)")
        verdict(*g.verify_points_host(h_proof.data_ptr(), h_comm.data_ptr(), h_z.data_ptr(), h_y.data_ptr(), n, rank * n, world * n), "group_verify_proof_batch(host)")

    def step_resident(record):
        if paz:
            return step_resident_paz(record)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if record else None
        if record: ev[0].record()
        check(s.blob_to_commitment_device(d_blobs.data_ptr(), n, d_comm.data_ptr(), d_status.data_ptr()), "blob_to_commitment")
        if record: ev[1].record()
        check(s.blob_proof_device(d_blobs.data_ptr(), d_comm.data_ptr(), n, d_proof.data_ptr(), d_status.data_ptr()), "compute_blob_proof")
        if record: ev[2].record()
        # shard partial + the ONLY data-path collective (ncclAllGather of 336 bytes per rank over NVLink) + pairing: one C-ABI call
        verdict(*g.verify_device([d_blobs.data_ptr()], [d_comm.data_ptr()], [d_proof.data_ptr()], [n], rank * n, world * n, [d_status.data_ptr()]),
                "group_verify_blob_proof_batch_device")
        if record: ev[3].record(); stage_ev.append(ev)

    def step_e2e():
        if paz:
            return step_e2e_paz()
        if args.e2e_api == "three_calls":
            check(s.blob_to_commitment_host(h_blobs.data_ptr(), n, h_comm.data_ptr()), "blob_to_commitment(host)")
            check(s.blob_proof_host(h_blobs.data_ptr(), h_comm.data_ptr(), n, h_proof.data_ptr()), "compute_blob_proof(host)")
            verdict(*g.verify_host(h_blobs.data_ptr(), h_comm.data_ptr(), h_proof.data_ptr(), n, rank * n, world * n), "group_verify_blob_proof_batch(host)")
        else:       # the same three pipelines over ONE upload of the blobs per step
            verdict(*g.commit_prove_verify_host(h_blobs.data_ptr(), n, rank * n, world * n, h_comm.data_ptr(), h_proof.data_ptr()), "group_commit_prove_verify(host)")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, *a):
        for _ in range(warmup):
            fn(*([False] * len(a)))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn(*a)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(local)
    # ---- resident arm
    for _ in range(args.warmup):
        step_resident(False)
    barrier()
    s.set_profiling(True)
    l0 = s.info()["kernel_launches"]
    if rank == 0: sampler.start()
    ms_res = timed(step_resident, args.steps, 0, True)
    clocks = sampler.stop() if rank == 0 else None
    launches = s.info()["kernel_launches"] - l0
    prof = s.msm_profile()
    s.set_profiling(False)
    # ---- end-to-end arm (host buffers)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    ms_e2e = timed(step_e2e, e2e_steps, 1)
    per_item = None
    if paz:
        # one wrong y on rank 0 must flip the RLC verdict on every rank; then PER-ITEM verdicts from the device pairing (kb200_verify_proof_many),
        # timed on their own: the wrong item, and only it, must come out false
        d_bad = d_y.clone()
        if rank == 0:
            d_bad[31] ^= 1
        verdict(*g.verify_points_device([d_proof.data_ptr()], [d_comm.data_ptr()], [d_z.data_ptr()], [d_bad.data_ptr()], [n], rank * n, world * n, [d_status.data_ptr()]),
                "group_verify_proof_batch_device(one wrong y)", want=False)
        h_bad = d_bad.cpu().numpy().tobytes()
        m = min(n, args.per_item)
        hp, hc, hz = bytes(h_proof.numpy()[: 48 * m]), bytes(h_comm.numpy()[: 48 * m]), bytes(h_z.numpy()[: 32 * m])
        s.verify_proof_many(hp, hc, hz, h_bad[: 32 * m], min(m, 64))             # warm-up (first-launch local-memory allocation)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ok, st = s.verify_proof_many(hp, hc, hz, h_bad[: 32 * m], m)
        dt = time.perf_counter() - t0
        if any(st) or ok != [not (rank == 0 and k == 0) for k in range(m)]:
            raise SystemExit("bench.py: per-item verdicts of the device pairing are wrong")
        per_item = {"items": m, "seconds": dt, "verdicts_per_s_per_gpu": m / dt, "api": "kb200_verify_proof_many (host buffers in, verdicts out)"}
        if not (torch.equal(h_proof, d_proof.cpu()) and torch.equal(h_y, d_y.cpu())):
            raise SystemExit("bench.py: host-buffer and resident outputs differ")
    else:
        # config 4's second run: one corrupted proof (rank 0, item 1 gets item 0's proof) must flip the verdict of the
        # whole sharded batch on every rank.  Untimed.
        d_bad = d_proof.clone()
        if rank == 0 and n > 1:
            d_bad[48:96] = d_proof[0:48]
        verdict(*g.verify_device([d_blobs.data_ptr()], [d_comm.data_ptr()], [d_bad.data_ptr()], [n], rank * n, world * n, [d_status.data_ptr()]),
                "group_verify_blob_proof_batch_device(one corrupted proof)", want=False)
        # parity guard inside the bench: the host-buffer path and the resident path must agree byte for byte
        if not (torch.equal(h_comm, d_comm.cpu()) and torch.equal(h_proof, d_proof.cpu())):
            raise SystemExit("bench.py: host-buffer and resident outputs differ")
    ginfo = g.info()
    g.close()
    return dict(rank=rank, world=world, n=n, s=s, ginfo=ginfo, per_item=per_item, h_z=(h_z if paz else None), h_y=(h_y if paz else None), ms_res=ms_res, ms_e2e=ms_e2e, e2e_steps=e2e_steps, clocks=clocks, launches=launches, prof=prof,
                stage_ev=stage_ev, info=info0, peak=(peak_wide, peak_fp), h_blobs=h_blobs, h_comm=h_comm, h_proof=h_proof)


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


_REAL_STDOUT = None


def keep_stdout_clean():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to stdout when
    NCCL_DEBUG=VERSION is set in the environment): point fd 1 at stderr for the rest of the run and keep the real stdout aside."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--blobs", type=int, default=4096, help="blobs per GPU per step (BASELINE configs[2]: 4096)")
    ap.add_argument("--workload", default="commit_prove_verify", choices=["commit_prove_verify", "proof_at_z"],
                    help="proof_at_z = BASELINE configs[4]: compute_kzg_proof at random points + verify_kzg_proof (RLC batch; per-item device pairing reported beside it)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="strong: --global-blobs are cut over the GPUs (north_star's scaling target)")
    ap.add_argument("--global-blobs", type=int, default=16384, help="batch size of --scaling strong (BASELINE configs[3]: 16384)")
    ap.add_argument("--window-bits", type=int, default=int(os.environ.get("KB200_WINDOW_BITS", "16")),
                    help="fixed-base table window width: 15 = 17 windows, 107 GiB table; 16 = GLV form, 2 x 8 windows on a 100 GiB table; library default 13 = 28.6 GiB")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-api", default="fused", choices=["fused", "three_calls"],
                    help="end-to-end arm: kb200_group_commit_prove_verify (one upload of the blobs per step) or the three separate host-buffer calls (three uploads)")
    ap.add_argument("--per-item", type=int, default=8192, help="proof_at_z: items per GPU checked with the device pairing (per-item verdicts)")
    ap.add_argument("--ref-blobs", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-verify-max", type=int, default=128, help="largest batch of the CPU baseline's batch-verify sweep (benches/kzg.rs:14: 1..128)")
    ap.add_argument("--no-cpu-breadth", action="store_true", help="skip the 1-core figure and the batch-verify sizes 1..128 of the CPU baseline")
    args = ap.parse_args()
    keep_stdout_clean()
    if args.impl == "reference":
        return run_reference(args)
    r = run_engine(args)
    rank, world, n, s = r["rank"], r["world"], r["n"], r["s"]
    if rank != 0:
        # leave together with rank 0 (it still formats and prints the line): no rank tears NCCL down while another is alive
        import torch.distributed as dist
        s.close()
        dist.barrier()
        dist.destroy_process_group()
        return
    import torch
    import kateth_b200 as K
    K_RECORD = K.RECORD_BYTES
    total = n * world
    ms_step = r["ms_res"] / args.steps
    value = total / (ms_step * 1e-3)
    ms_e2e = r["ms_e2e"] / r["e2e_steps"]
    # per-stage device times (rank 0), averaged over the timed steps
    paz = args.workload == "proof_at_z"
    nst = 2 if paz else 3
    st = [[ev[i].elapsed_time(ev[i + 1]) for i in range(nst)] for ev in r["stage_ev"]]
    stage = [sum(x[i] for x in st) / len(st) for i in range(nst)]
    stage_names = ["prove_at_z", "verify_rlc"] if paz else ["commit", "prove", "verify"]
    # roofline of the dominant kernel (k_msm): SURVEY 8(d) yardstick MAC32 per MSM x MSMs per launch / CUDA-event launch duration
    prof, info = r["prof"], r["info"]
    peak_chain, peak_fp = r["peak"]
    peak_wide = max(peak_chain, peak_fp)      # both streams are IMAD.WIDE-bound; take the better of the two as the denominator
    msm_ms = prof["ms"] / max(1, prof["launches"])
    msms_per_launch = prof["msms"] / max(1, prof["launches"])
    achieved = MAC32_PER_MSM * msms_per_launch / (msm_ms * 1e-3) / 1e12 if msm_ms > 0 else 0.0
    peaks = load_peaks()
    hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
    info1 = s.info()
    W = info["windows"] * (info.get("glv", 0) + 1)      # gathered points per scalar (GLV tables: two halves share the windows)
    ba = bool(info1.get("last_msm_batch_affine", 0))
    if ba:
        # k_msm_ba (msm_ba.cuh): per thread 31W affine additions (6 Fp products each: 1 prefix, 2 suffix, 1 slope, 1 square,
        # 1 product), W mixed additions (10), 5 inversions (1 product + ~3120 MAC32 of 30-bit linear combinations each) and
        # its share of the CTA tree reduction (127 full additions of 14 products per 128 threads); 128 threads per MSM.
        fp_products = 128.0 * (6 * 31 * W + 10 * W + 5 * 1 + 14)
        mac_per_msm = fp_products * 300.0 + 128.0 * 5 * 3120.0
        gather_per_msm = 4096.0 * W * (64 + 96)          # every table point is gathered twice: x (2 sectors), then x and y
        kernel = "k_msm_ba (+ k_msm_recode)"
        executed_note = ("executed = MAC32 actually issued by the batch-affine MSM: 128 threads x (6 x 31W affine additions + 10 x W mixed "
                         "additions + 5 binary-GCD inversions + tree reduction), W = %d windows; achieved = SURVEY 8d yardstick" % W)
    else:
        mac_per_msm = 4096.0 * W * 3000.0
        gather_per_msm = 4096.0 * W * 96
        kernel = "k_msm"
        executed_note = "executed = 4096 x windows x 3000 MAC32 (XYZZ mixed additions actually issued) per MSM; achieved = SURVEY 8d yardstick"
    gather_bytes = gather_per_msm * msms_per_launch
    traffic = None
    try:        # dram__bytes_read.sum + dram__bytes_write.sum of the committed ncu --set full capture, scaled to this launch
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_k_msm_ba_traffic.json")))
        if ba:
            traffic = tr["dram_bytes_per_msm"] * msms_per_launch
    except Exception:
        pass
    executed = mac_per_msm * msms_per_launch / (msm_ms * 1e-3) / 1e12 if msm_ms > 0 else 0.0
    # roofline = multiplier-pipe utilisation: MAC32 the kernel actually issues per launch / launch time / measured IMAD.WIDE rate.
    # The SURVEY 8d yardstick (bucket Pippenger, 2.80e8 MAC32 per MSM) is reported beside it as an ALGORITHMIC ratio: the fixed-base
    # table does the yardstick's job with fewer multiplications (bought with HBM), which is not pipe utilisation.
    roofline = {"bound": "imad", "achieved": executed, "peak": peak_wide, "unit": "TMAC32/s", "frac": executed / peak_wide if peak_wide else None,
                "traffic": traffic, "traffic_note": "DRAM bytes per launch: ncu capture of a 1184-MSM launch (profiles/r02_k_msm_ba_traffic.json) scaled by MSMs per launch; table gathers + per-CTA scratch of the affine rounds",
                "kernel": kernel, "launch_ms": msm_ms, "msms_per_launch": msms_per_launch, "share_of_step": prof["ms"] / r["ms_res"],
                "executed_mac32_per_msm": mac_per_msm,
                "peak_source": "K0 measured in this run: max(IMAD.WIDE.U32.X carry chains %.2f, Fp-product chain %.2f) TMAC32/s; IMAD.WIDE issues at half the 32-bit IMAD rate on sm_100a; no integer peak in MEASURED_PEAKS.json" % (peak_chain, peak_fp),
                "achieved_note": executed_note.replace("; achieved = SURVEY 8d yardstick", ""),
                "yardstick": {"mac32_per_msm": MAC32_PER_MSM, "equivalent_tmac32_per_s": achieved, "algorithmic_speedup": MAC32_PER_MSM / mac_per_msm,
                              "note": "SURVEY 8d fixed yardstick: signed-digit c=13 bucket Pippenger, 10 Fp products per mixed addition, 300 MAC32 per Fp product; "
                                      "equivalent = yardstick MAC32 x MSMs / launch time -- NOT a pipe utilisation (it exceeds the peak because the table removes multiplications)"},
                "hbm": {"achieved": gather_bytes / (msm_ms * 1e-3) / 1e9 if msm_ms > 0 else 0.0, "peak": hbm_peak, "unit": "GB/s",
                        "frac": (gather_bytes / (msm_ms * 1e-3) / 1e9 / hbm_peak) if msm_ms > 0 else None,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
                        "algorithmic": "table sectors gathered per MSM (4096 x %d windows%s); scratch traffic of the affine rounds not counted" % (W, ", twice" if ba else "")}}
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as O
        threads = host_threads()
        m = min(n, max(threads, 8))
        # the sample strides across the whole batch (first wave of the persistent MSM grid, later waves, the last item), and the
        # bench doubles as a parity check: the oracle's bytes for the sample must equal the engine's
        idx = sorted(set(int(round(k * (n - 1) / max(1, m - 1))) for k in range(m)))
        hb = r["h_blobs"].numpy()
        sample = b"".join(bytes(hb[i * BLOB:(i + 1) * BLOB]) for i in idx)
        hc, hp = r["h_comm"].numpy(), r["h_proof"].numpy()
        caveat = ("portable-C restatement of kateth's algorithm (oracle/), not blst's assembly with its thread-pooled Pippenger: kateth itself would be "
                  "faster per core; Rust and blst are absent from this image, so it is the only CPU baseline obtainable here")
        if paz:
            o = cpu_leg.setup = getattr(cpu_leg, "setup", None) or O.Setup.load_json(SETUP_JSON)
            hz, hy = r["h_z"].numpy(), r["h_y"].numpy()
            t0 = time.perf_counter()
            for k, i in enumerate(idx[:4]):           # Setup::proof + Setup::verify_proof per item, one core (oracle.proof has no threaded leg)
                pr, y = o.proof(sample[k * BLOB:(k + 1) * BLOB], bytes(hz[32 * i:32 * i + 32]))
                if pr != bytes(hp[48 * i:48 * i + 48]) or y != bytes(hy[32 * i:32 * i + 32]):
                    raise SystemExit("bench.py: engine (proof, y) differs from the oracle on the CPU-baseline sample")
                if not o.verify_proof(pr, bytes(hc[48 * i:48 * i + 48]), bytes(hz[32 * i:32 * i + 32]), y):
                    raise SystemExit("bench.py: oracle rejects its own proof")
            dt = time.perf_counter() - t0
            cpu_baseline = {"value": min(4, len(idx)) / dt, "unit": "blobs/s", "cores": 1, "kind": "port",
                            "sample": "%d blobs strided across the step's batch, Setup::proof + Setup::verify_proof (one pairing each) on one core, %.2f s; outputs equal the engine's" % (min(4, len(idx)), dt),
                            "caveat": caveat}
        else:
            a, b, comm, proofs = cpu_leg(sample, len(idx), threads)
            if comm != b"".join(bytes(hc[48 * i:48 * i + 48]) for i in idx) or proofs != b"".join(bytes(hp[48 * i:48 * i + 48]) for i in idx):
                raise SystemExit("bench.py: engine output differs from the oracle on the CPU-baseline sample")
            cpu_baseline = {"value": len(idx) / (a + b), "unit": "blobs/s", "cores": threads, "kind": "port",
                            "sample": "%d blobs strided across the step's batch (items %s), %d threads (commit+prove %.2f s, batch-verify %.2f s); outputs equal the engine's"
                                      % (len(idx), ",".join(str(i) for i in idx), threads, a, b),
                            "caveat": caveat}
            if not args.no_cpu_breadth:
                # the 1-core figure, and kateth's own bench shape -- one verify_blob_kzg_proof_batch call per size 1..128
                # (benches/kzg.rs:14,45-64; kateth's batch verification is single-threaded: naive lincombs + one pairing)
                a1, b1, _, _ = cpu_leg(sample[: 2 * BLOB], 2, 1)
                cpu_baseline["one_core"] = {"value": 2 / (a1 + b1), "unit": "blobs/s", "cores": 1, "sample": "2 blobs: commit+prove %.2f s, batch-verify %.2f s" % (a1, b1),
                                            "commit_prove_blobs_per_s": 2 / a1, "batch_verify_blobs_per_s": 2 / b1}
                o = cpu_leg.setup
                sizes, cur = {}, 1
                nb = min(n, args.cpu_verify_max)
                vb, vc, vp = bytes(hb[: nb * BLOB]), bytes(hc[: 48 * nb]), bytes(hp[: 48 * nb])     # the engine's own commitments / proofs as inputs
                while cur <= nb:
                    t0 = time.perf_counter()
                    ok = o.verify_many(vb[: cur * BLOB], vc[: 48 * cur], vp[: 48 * cur], cur, 1)
                    sizes[str(cur)] = cur / (time.perf_counter() - t0)
                    if not ok:
                        raise SystemExit("bench.py: the oracle rejects the engine's proofs (batch of %d)" % cur)
                    cur *= 2
                cpu_baseline["batch_verify_by_size_one_core_blobs_per_s"] = sizes
    e2e_bytes = ({"h2d_bytes_per_step": n * BLOB + n * 32 + n * (48 + 48 + 32 + 32), "d2h_bytes_per_step": n * (48 + 32) + 8 * n + 200} if paz else
                 {"h2d_bytes_per_step": 3 * n * BLOB + 3 * n * 48, "d2h_bytes_per_step": 2 * n * 48 + 200 + 8 * n} if args.e2e_api == "three_calls" else
                 {"h2d_bytes_per_step": n * BLOB, "d2h_bytes_per_step": 2 * n * 48 + 216 + 24 * n})
    line = {"metric": METRIC_PAZ if paz else METRIC, "value": value, "unit": "blobs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u32 limbs (Fp 12x32, Fr 8x32 Montgomery, IMAD.WIDE)",
            "data": "synthetic canonical field elements, uniform in [0, r)",
            "config": {"workload": ("compute_kzg_proof at random points + verify_kzg_proof (RLC batch over all GPUs) over %d blobs per GPU (BASELINE configs[4])" % n) if paz else
                                   ("commit + compute_blob_kzg_proof + verify_blob_kzg_proof_batch over %d blobs per GPU (BASELINE configs[2] size; configs[3] sharding)" % n),
                       "blobs_per_gpu": n, "global_blobs": total, "window_bits": info["window_bits"], "windows": info["windows"], "glv": bool(info.get("glv", 0)), "points_per_scalar": W, "msm_shape": "batch-affine (k_msm_ba)" if ba else "xyzz (k_msm)",
                       "table_gb": round(info["table_bytes"] / 2 ** 30, 2), "parallelism": "blob shards x%d, replicated tables, 1 ncclAllGather of %d B per rank inside the C ABI (kb200_group_*, NCCL %s)" % (world, K_RECORD, r["ginfo"]["nccl_version"] or "not needed"),
                       "l2": "inputs (%d MiB of blobs + multi-GB table) exceed the 126 MB L2; no flush needed" % (n * BLOB >> 20)},
            "stages_ms": {k: v for k, v in zip(stage_names, stage)},
            "stage_blobs_per_s_per_gpu": {k: n / v * 1e3 for k, v in zip(stage_names, stage)},
            "e2e": {"value": total / (ms_e2e * 1e-3), "unit": "blobs/s", **e2e_bytes,
                    "ms_per_step": ms_e2e, "steps": r["e2e_steps"],
                    "api": "kb200_compute_proof + kb200_group_verify_proof_batch (pinned host buffers)" if paz else
                           ("kb200_blob_to_commitment + kb200_compute_blob_proof + kb200_group_verify_blob_proof_batch (pinned host buffers, three uploads)" if args.e2e_api == "three_calls"
                            else "kb200_group_commit_prove_verify: commit + blob proof + batch verification, each complete, over one upload of the blobs (pinned host buffers)")},
            "per_item_verdicts": r["per_item"],
            "gpu_launches": int(r["launches"]), "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": r["clocks"],
            "setup_seconds": info["setup_seconds"],
            "checks": {"valid_batch_verdict": True, "one_corrupted_item_verdict": False, "host_buffer_outputs_equal_resident": True,
                       "oracle_outputs_equal_engine_on_cpu_sample": True if cpu_baseline is not None else None}}      # None: CPU leg not run (N > 1 or --no-cpu-baseline)
    emit(line)
    s.close()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
