#!/usr/bin/env python
"""bench.py -- LipVQ-VAE tokenizer hot path on B200: action chunks tokenized / s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A "step" = one encode+quantize+decode pass (LLFQVAE_V4.forward, v5:70-84) over one batch of
synthetic action chunks (BASELINE.json config 3's batch: 65 536 chunks = 655 360 rows x 12-D,
D=208, K=1024, fp32, reference-distribution random-init weights) PER GPU (weak scaling; rows are
independent so the forward has no data-path collective).  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

F_DIM, D_LAT, K_CODES, H_DIM = 12, 208, 1024, 128
CHUNKS, CHUNK_LEN = 65536, 10
FLOP_PER_ROW_SEARCH = 2 * K_CODES * D_LAT
ZB_BYTES = 2 * 16 * ((D_LAT + 3 + 15) // 16)                  # fp16 row operand [z - mu | 1 1 1], whole MMA K steps
WORKLOAD = ("tokenizer encode+quantize+decode, %d chunks x %d rows x %d-D per GPU, D=%d, K=%d, fp32"
            % (CHUNKS, CHUNK_LEN, F_DIM, D_LAT, K_CODES))


_REAL_STDOUT = None


def _emit(obj):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(obj) + "\n")
    out.flush()


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _make_params(seed=0):
    from oracle import lipvq_oracle as orc     # input generator only (numpy RNG, same as the tests)
    return orc.init_params(F_DIM, D_LAT, K_CODES, H_DIM, seed=seed)


# ------------------------------------------------------------------------------------------ CPU
def cpu_reference_sample(rows, threads, search="reference"):
    """Time the oracle port of the reference forward on `rows` rows with `threads` host threads.
    Row blocks run in a thread pool (numpy releases the GIL inside ufuncs/BLAS)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import lipvq_oracle as orc
    p = _make_params()
    x = np.random.default_rng(1).standard_normal((rows, F_DIM)).astype(np.float32)
    blk = max(64, (rows + threads - 1) // threads)
    parts = [x[s:s + blk] for s in range(0, rows, blk)]

    def run(xb):
        return orc.forward(p, xb, search=search)["loss"]

    t0 = time.perf_counter()
    if threads > 1:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(run, parts))
    else:
        [run(xb) for xb in parts]
    return time.perf_counter() - t0


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample_chunks = 256                                       # BASELINE config 1's batch
    rows = sample_chunks * CHUNK_LEN
    for _ in range(min(args.warmup, 1)):
        cpu_reference_sample(rows, cores)
    ts = [cpu_reference_sample(rows, cores) for _ in range(args.steps)]
    ms = 1e3 * sum(ts) / len(ts)
    val = sample_chunks / (ms / 1e3)
    sample = ("%d chunks (%d rows) per step, numpy port of the reference's broadcast-form forward "
              "(v5:37-48 literally), %d threads over row blocks" % (sample_chunks, rows, cores))
    # what the reference's own entry point does: torch.set_num_threads(1) (robomimic/scripts/train.py:57)
    t1 = min(cpu_reference_sample(rows, 1) for _ in range(2))
    _emit({
        "impl": "reference", "metric": "action chunks tokenized/sec (encode+quantize+decode)", "value": val,
        "unit": "chunks/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD + " -- CPU arm: numpy PORT of the reference forward timed on a %d-chunk SAMPLE "
                               "of that batch and extrapolated per chunk (rows are independent); the reference itself "
                               "is a Python/torch module that cannot travel to this box" % sample_chunks,
                   "sample": sample},
        "cpu_baseline": {"value": val, "unit": "chunks/s", "cores": cores, "kind": "port", "sample": sample,
                         "one_thread_value": sample_chunks / t1,
                         "one_thread_note": "1 thread, as robomimic/scripts/train.py:57 pins torch"},
        "e2e": {"value": val, "unit": "chunks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


# ------------------------------------------------------------------------------------------ GPU
def _time_events(fn, iters, torch, stream=None):
    """average ms per call, CUDA events on the launching (current) stream."""
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s.record()
    for i in range(iters):
        fn(i)
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def run_ours(args, rank, world, local_rank):
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from lipvq_vae_b200 import LLFQVAE_V4, SEARCH_EXACT, SEARCH_TENSOR, _native as N

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # one process per GPU: stage the pinned host batches on the GPU's own NUMA node (matters for `e2e` at N > 1)
    from lipvq_vae_b200 import bind_host_thread_to_gpu_node
    numa_cpus = bind_host_thread_to_gpu_node(local_rank) if world > 1 else None
    lib = N.lib()
    N.check(lib.lipvq_check_device(), "lipvq_check_device")
    peaks, peak_src = _peaks()

    rows = CHUNKS * CHUNK_LEN
    p = _make_params()
    m = LLFQVAE_V4(F_DIM, D_LAT, num_codes=K_CODES, hidden_dim=H_DIM).to(dev)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})
    m.search_mode = SEARCH_EXACT if args.search == "exact" else SEARCH_TENSOR
    m.eval()

    # rotating inputs: NBUF x 31.5 MB > the 126 MB L2, so no step re-reads a cached input
    NBUF = 6
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    host = [torch.randn(rows, F_DIM, generator=g).pin_memory() for _ in range(NBUF)]
    xs = [h.to(dev) for h in host]
    loss_host = torch.zeros(4).pin_memory()
    idx_host = torch.zeros(rows, dtype=torch.int64).pin_memory()

    def step(i):
        with torch.no_grad():
            return m(xs[i % NBUF])

    def step_e2e(i):
        xd = xs[i % NBUF]
        xd.copy_(host[i % NBUF], non_blocking=True)            # H2D of this step's inputs
        with torch.no_grad():
            z, loss = m(xd)
        loss_host.copy_(m.last_loss_terms, non_blocking=True)  # D2H of the step's result
        idx_host.copy_(m.last_indices, non_blocking=True)      # ... and its tokens
        torch.cuda.current_stream().synchronize()

    def barrier():
        if world > 1:
            dist.barrier()

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    l0 = lib.lipvq_launch_count()
    ms = _time_events(step, args.steps, torch)
    launches = lib.lipvq_launch_count() - l0
    barrier()
    clk = clocks.stop() if rank == 0 else None

    for i in range(min(args.warmup, 3)):
        step_e2e(i)
    barrier()
    ms_e2e_serial = _time_events(step_e2e, args.steps, torch)
    barrier()
    # the public host-batch API (lipvq_vae_b200.HostTokenizer): every step still uploads its own pinned-host
    # batch and brings its tokens + loss terms back, but batch i+1's upload and batch i-1's download run on the
    # copy engines while batch i's kernels run.  All K uploads / K forwards / K downloads are inside the region.
    from lipvq_vae_b200 import HostTokenizer
    DEPTH = int(os.environ.get("LIPVQ_E2E_DEPTH", "2"))
    ht = HostTokenizer(m, rows, depth=DEPTH)

    def e2e_loop(k):
        sink = 0.0
        for i in range(k):
            ht.submit(host[i % NBUF])
            if i >= DEPTH - 1:
                sink += float(ht.collect()[1][0])          # host reads the loss of the oldest batch in flight
        ht.drain_to()
        return sink

    def e2e_flush():
        while ht.pending:
            ht.collect()
    e2e_loop(min(args.warmup, 3)); e2e_flush(); torch.cuda.synchronize()
    barrier()
    s_ev, e_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s_ev.record()
    e2e_loop(args.steps)
    e_ev.record()
    e2e_flush()
    torch.cuda.synchronize()
    ms_e2e = s_ev.elapsed_time(e_ev) / args.steps
    barrier()

    if world > 1:
        t = torch.tensor([ms, ms_e2e, ms_e2e_serial], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e, ms_e2e_serial = t.tolist()
    value = world * CHUNKS / (ms / 1e3)
    e2e = world * CHUNKS / (ms_e2e / 1e3)

    # ---- BASELINE config 3: the data-parallel training step on ALL ranks -- forward + backward of the tokenizer loss
    #      on this rank's 65 536 chunks, ONE NCCL all-reduce of the flat gradient bucket (+ the loss scalar), ONE fused
    #      AdamW launch over the 14 tensors (icl.py:886-889, 913-914, 968-970).  Weak scaling like the forward.
    import copy
    from lipvq_vae_b200 import TokenizerTrainer
    m_tr = copy.deepcopy(m).train()
    trainer = TokenizerTrainer(m_tr)
    n_total = world * rows
    tr_steps = max(5, args.steps // 2)

    def train_step(i):
        trainer.step(xs[i % NBUF], n_total)
    for i in range(max(3, args.warmup)):
        train_step(i)
    torch.cuda.synchronize()
    barrier()
    lt0 = lib.lipvq_launch_count()
    ms_train = _time_events(train_step, tr_steps, torch)
    train_launches = (lib.lipvq_launch_count() - lt0) / tr_steps
    barrier()
    if world > 1:
        t = torch.tensor([ms_train], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_train = float(t.item())
    train = {"workload": "training step: forward + backward (14 gradients) + flat-bucket all-reduce + fused AdamW, "
                         "%d chunks per GPU" % CHUNKS,
             "n_gpus": world, "ms_per_step": ms_train, "chunks_per_s": world * CHUNKS / (ms_train / 1e3),
             "steps": tr_steps, "own_kernel_launches_per_step": train_launches,
             "allreduce_bytes": int(trainer.dp._bucket.flat.numel() * 4), "timing": "CUDA events, max over ranks"}
    # per-kernel split of the training step (library CUDA events around each launch, after the timed region).  EVERY
    # rank runs these three steps -- each contains the gradient all-reduce -- and only rank 0 keeps the report.
    # (with the decoder-side and encoder-side backward on ONE stream: on two streams the event pairs of concurrent
    #  kernels overlap and every duration is inflated)
    trainer.concurrent_backward = False
    rep = N.profile_kernels(train_step, 3)
    trainer.concurrent_backward = True
    torch.cuda.synchronize()
    barrier()
    if rank == 0:
        # the two backward GEMM families report against their rooflines: weight gradients 2*N*sum(Nout*Kin) flop as
        # 3 tf32 MMAs per product, input gradients likewise; both also move every saved activation once
        train["kernels_us_per_step"] = {k: 1e3 * v[1] / 3 for k, v in rep.items()}
        Hh, H1h, D, F = H_DIM, 64, D_LAT, F_DIM
        hbm_peak_for_train = _peaks()[0]["hbm_gbs"]
        # (the first layer's weight gradient -- Kin = 12 -- is the streaming fp32 kernel wgrad_narrow_kernel)
        wg_flop = 2.0 * rows * (D * Hh + Hh * H1h)
        wn_flop = 2.0 * rows * (H1h * F)
        dg_flop = 2.0 * rows * (D * Hh + Hh * H1h)
        wg_bytes = 4.0 * rows * (D + Hh + Hh + H1h)
        wn_bytes = 4.0 * rows * (H1h + F)
        dg_bytes = 4.0 * rows * (D + 2 * Hh + Hh + 2 * H1h)
        for kname, flop, byts in (("wgrad_umma_kernel", wg_flop, wg_bytes), ("wgrad_narrow_kernel", wn_flop, wn_bytes),
                                  ("dgrad_umma_kernel", dg_flop, dg_bytes)):
            us = train["kernels_us_per_step"].get(kname)
            if us:
                train.setdefault("backward_rooflines", []).append({
                    "kernel": kname, "us_per_step": us, "bound": "hbm", "achieved": byts / (us * 1e-6) / 1e9,
                    "peak": hbm_peak_for_train, "unit": "GB/s", "frac": byts / (us * 1e-6) / 1e9 / hbm_peak_for_train,
                    "tensor_tflops_fp32_equivalent": flop / (us * 1e-6) / 1e12,
                    "algorithmic": "reads of dY and the saved activations + writes of the activation gradients, fp32, once each"})
    del trainer, m_tr
    if world > 1:
        # A/B of the collective: the split schedule (first half of the bucket all-reduced UNDER the encoder backward)
        # against the default above (ONE all-reduce after the whole backward) -- measured no better, see dist.py
        m_tr = copy.deepcopy(m).train()
        trainer = TokenizerTrainer(m_tr, overlap=True)
        for i in range(3):
            train_step(i)
        torch.cuda.synchronize()
        barrier()
        ms_serial = _time_events(train_step, tr_steps, torch)
        barrier()
        t = torch.tensor([ms_serial], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        train["ms_per_step_split_allreduce_under_backward"] = float(t.item())
        del trainer, m_tr
    # A/B: decoder-side and encoder-side backward on ONE stream (the default runs them on two)
    m_tr = copy.deepcopy(m).train()
    trainer = TokenizerTrainer(m_tr, concurrent_backward=False)
    for i in range(3):
        train_step(i)
    torch.cuda.synchronize()
    barrier()
    ms_cc = _time_events(train_step, tr_steps, torch)
    barrier()
    if world > 1:
        t = torch.tensor([ms_cc], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_cc = float(t.item())
    train["ms_per_step_serial_backward"] = ms_cc
    del trainer, m_tr
    train["collective"] = ("ONE NCCL all-reduce per step over the flat gradient bucket (14 tensors + the loss scalar), after "
                           "the backward; gradients are written by the kernels straight into the bucket (no pack/unpack "
                           "copies).  The split schedule that puts 80 % of the bucket under the encoder backward is the "
                           "A/B leg ms_per_step_split_allreduce_under_backward: no gain (the collective is ~40 us alone "
                           "and competes for SMs with the persistent one-CTA-per-SM backward kernels)")
    # EMA variant of the same step (north_star item 3; extension, not in the reference): the codebook follows the EMA
    # of its clusters -- per-code counts/sums by the scatter-add kernel, all-reduced in one more collective
    m_tr = copy.deepcopy(m).train()
    trainer = TokenizerTrainer(m_tr, ema_codebook=True)
    for i in range(2):
        train_step(i)
    torch.cuda.synchronize()
    barrier()
    ms_ema = _time_events(train_step, tr_steps, torch)
    barrier()
    if world > 1:
        t = torch.tensor([ms_ema], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_ema = float(t.item())
    train["ema_codebook_variant"] = {"ms_per_step": ms_ema, "chunks_per_s": world * CHUNKS / (ms_ema / 1e3),
                                     "extra": "scatter-add of z_e per code + all-reduce of counts[K], sums[K,D] + "
                                              "lipvq_ema_update (parity unpinned: no reference counterpart)"}
    del trainer, m_tr
    torch.cuda.empty_cache()
    skip = set(filter(None, args.skip.split(",")))
    # ---- BASELINE configs 4 and 5 run on EVERY rank (weak scaling, max over ranks inside the legs)
    c4 = None if "config4" in skip else _guard(config4_leg, torch, N, lib, dev, rank, world, peaks, args.steps)
    torch.cuda.empty_cache()
    c5 = None if "config5" in skip else _guard(config5_leg, torch, dev, rank, world, args.steps)
    torch.cuda.empty_cache()

    if rank != 0:
        return
    parity = None if "parity" in skip else parity_leg(torch, m, p, xs[0])      # raises SystemExit on a mismatch
    latency = None if "latency" in skip else _guard(latency_leg, torch, N, dev, p)
    vqc = None if "vqvae" in skip else _guard(vqvae_classic_leg, torch, dev, rows, args.steps)
    # ---- per-kernel durations INSIDE the step: the library brackets each of its launches with CUDA events on
    #      the launching stream (lipvq_profile_begin/end); same workload, same rotating inputs
    prof_iters = max(5, args.steps)
    prof = N.profile_kernels(step, prof_iters)
    kern_us = {k: 1e3 * ms_tot / prof_iters for k, (cnt, ms_tot) in prof.items()}     # us per step
    kern_launches = {k: cnt / prof_iters for k, (cnt, ms_tot) in prof.items()}
    stats_now = m.last_search_stats.cpu()
    sustained = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])

    def _roof(name, bound, algo, peak, unit, note):
        us = kern_us.get(name)
        if not us:
            return None
        ach = algo / (us * 1e-6) / (1e12 if unit == "TFLOP/s" else 1e9)
        return {"bound": bound, "kernel": name, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
                "traffic": None, "us_per_launch": us / max(kern_launches.get(name, 1.0), 1.0),
                "share_of_step": us / (ms * 1e3), "algorithmic": note}
    rooflines = [r for r in (
        _roof("vq_tc_kernel", "tensor", rows * FLOP_PER_ROW_SEARCH, sustained, "TFLOP/s",
              "2*K*D flop/row = %d (fp16 operands, fp32 accumulate; SURVEY 8d)" % FLOP_PER_ROW_SEARCH),
        # the encoder's binding roof at the optimum is HBM: its stores (z_e + the search operand) take 133 us at the
        # measured copy bandwidth, its 3-term fp16 split products 98 us at the measured tensor rate
        _roof("enc16_fused_kernel", "hbm", rows * (F_DIM * 4.0 + D_LAT * 4 + ZB_BYTES), peaks["hbm_gbs"], "GB/s",
              "read x (F*4 B) + write z_e (D*4 B) + write the search's fp16 row operand (%d B) per row; its GEMMs are "
              "71168 fp32-equivalent flop/row run as 3 kind::f16 MMAs per product (tensor floor 98 us < HBM floor "
              "133 us), see `tensor_tflops_fp32_equivalent`" % ZB_BYTES),
        _roof("quantize_emit_kernel", "hbm", rows * (2.0 * D_LAT * 4 + 8 + 2 * F_DIM * 4), peaks["hbm_gbs"], "GB/s",
              "read z_e + write z_q (2*D*4 B/row) + idx 8 B + x read and x_recon write (2*F*4 B)"),
    ) if r]
    # dram bytes per launch of the same kernels on the same workload, from the committed ncu --set full capture
    try:
        _traffic = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))["kernels"]
    except Exception:
        _traffic = {}
    for r in rooflines:
        t_ = _traffic.get(r["kernel"])
        r["traffic"] = t_["total_bytes"] if t_ else None
        r["traffic_source"] = "profiles/r02_traffic.json (ncu dram__bytes_read.sum + dram__bytes_write.sum)" if t_ else None
    for r in rooflines:
        r["peak_source"] = peak_src + (" bf16 sustained (kernel timed inside the step)" if r["bound"] == "tensor"
                                       else " HBM copy bandwidth")
    for r in rooflines:
        if r["kernel"] == "enc16_fused_kernel":
            r["tensor_tflops_fp32_equivalent"] = rows * 2.0 * (F_DIM * 64 + 64 * H_DIM + H_DIM * D_LAT) / (
                r["us_per_launch"] * 1e-6) / 1e12
            r["peak_source"] = peak_src + " HBM copy bandwidth"
    roofline = max(rooflines, key=lambda r: r["share_of_step"]) if rooflines else None
    if roofline is not None:
        roofline = dict(roofline)
        roofline["fallback_rows"] = int(stats_now[0].item())
        roofline["rescored_codes"] = int(stats_now[1].item())

    z_e = torch.empty(rows, D_LAT, device=dev)
    with torch.no_grad():
        z_e.copy_(m.encode(xs[0]))
    E = m.quantizer.codebook.detach()
    idx = torch.empty(rows, dtype=torch.int64, device=dev)
    def _ws_bytes(n, D, K):
        need = C.c_size_t(0)
        N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
        return need.value
    ws = torch.empty(max(_ws_bytes(rows, D_LAT, K_CODES), _ws_bytes(1 << 20, 64, 1024)) + 4096,
                     dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    mode = int(m.search_mode)
    stats = torch.zeros(4, dtype=torch.int32, device=dev)

    def vq(i, z=z_e, Ecb=E, n=rows, D=D_LAT, K=K_CODES):
        N.check(lib.lipvq_vq_argmin(z.data_ptr(), Ecb.data_ptr(), n, D, K, mode, idx.data_ptr(), None,
                                    stats.data_ptr(), ws.data_ptr(), ws.numel(), st))
    for i in range(3):
        vq(i)
    # ---- BASELINE config 2: the search kernel alone, N=2^20, D=64, K=1024, fp32
    n2, d2, k2 = 1 << 20, 64, 1024
    z2 = torch.rand(n2, d2, device=dev)
    E2 = torch.rand(k2, d2, device=dev)
    idx2 = torch.empty(n2, dtype=torch.int64, device=dev)

    def vq2(i):
        N.check(lib.lipvq_vq_argmin(z2.data_ptr(), E2.data_ptr(), n2, d2, k2, mode, idx2.data_ptr(), None,
                                    stats.data_ptr(), ws.data_ptr(), ws.numel(), st))
    for i in range(3):
        vq2(i)
    ms2 = _time_events(vq2, max(5, args.steps // 2), torch)
    prof2 = N.profile_kernels(vq2, 5)
    tf2 = n2 * 2.0 * k2 * d2 / (ms2 / 1e3) / 1e12
    c2 = {"workload": "VQ nearest-codeword kernel alone: N=2^20, D=64, K=1024, fp32", "ms": ms2,
          "tflops": tf2, "frac_of_bf16_peak": tf2 / peaks["bf16_tflops"],
          "algorithmic_gbs": (n2 * (4 * d2 + 8) + k2 * d2 * 4) / (ms2 / 1e3) / 1e9,
          "fallback_rows": int(stats[0].item()), "rescored_codes": int(stats[1].item()),
          "kernels_us": {k: 1e3 * v[1] / 5 for k, v in prof2.items()},
          "vq_tc_kernel_tflops": (n2 * 2.0 * k2 * d2 / (prof2["vq_tc_kernel"][1] / 5 / 1e3) / 1e12
                                  if "vq_tc_kernel" in prof2 else None)}
    # ---- HBM-bound gather+loss pass (k4/k5): 3 x N x D x 4 bytes
    zq = torch.empty_like(z_e)
    sse = torch.zeros(1, dtype=torch.float64, device=dev)

    def gl(i):
        N.check(lib.lipvq_gather_loss(z_e.data_ptr(), E.data_ptr(), idx.data_ptr(), rows, D_LAT, K_CODES,
                                      zq.data_ptr(), sse.data_ptr(), ws.data_ptr(), ws.numel(), st))
    vq(0)
    for i in range(3):
        gl(i)
    ms_gl = _time_events(gl, max(5, args.steps // 2), torch)
    gbs = (rows * D_LAT * 4 * 2 + rows * 8) / (ms_gl / 1e3) / 1e9
    hbm = {"kernel": "quantize_emit (gather + commitment-SSE mode, via lipvq_gather_loss)", "ms": ms_gl, "achieved_gbs": gbs, "peak_gbs": peaks["hbm_gbs"],
           "frac": gbs / peaks["hbm_gbs"], "bytes": "read z_e + write z_q + idx (codebook rows hit L2)"}

    cpu_baseline = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        sample_chunks = 256
        ts = [cpu_reference_sample(sample_chunks * CHUNK_LEN, cores) for _ in range(8)]     # ~5 s of CPU work
        t = min(ts)
        t_gemm = min(cpu_reference_sample(sample_chunks * CHUNK_LEN * 4, cores, search="gemm") for _ in range(4))
        cpu_baseline = {"value": sample_chunks / t, "unit": "chunks/s", "cores": cores, "kind": "port",
                        "sample": "%d chunks (%d rows): numpy port of the reference's broadcast-form forward "
                                  "(v5:37-48), %d threads over row blocks, best of 8 (mean in `mean_value`)" % (sample_chunks,
                                                                                       sample_chunks * 10, cores),
                        "mean_value": sample_chunks / (sum(ts) / len(ts)),
                        "gemm_form_value": 4 * sample_chunks / t_gemm,
                        "gemm_form_note": "same math with the search as ||e||^2-2zE^T in fp32 BLAS (SURVEY 6)"}
    out = {
        "metric": "action chunks tokenized/sec (encode+quantize+decode)", "value": value, "unit": "chunks/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "search": "exact" if mode == 0 else "tensor",
                   "outputs": "z_latent [N,D], token indices [N], reconstruction [N,F], loss terms",
                   "weights": "reference init distributions, numpy seed 0",
                   "l2": "%d rotating input buffers (%.0f MB) + %.1f GB of per-step intermediates exceed the 126 MB L2"
                         % (NBUF, NBUF * rows * F_DIM * 4 / 1e6, rows * (2 * D_LAT * 4 + 512 + 8 + F_DIM * 4) / 1e9)},
        "e2e": {"value": e2e, "unit": "chunks/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": rows * F_DIM * 4, "d2h_bytes_per_step": rows * 8 + 16,
                "api": "z_latent stays on the device (its consumer, the policy embedding, is there: obs_nets.py:1336, 2582); "
                       "HostTokenizer.submit(pinned host batch) / collect(): per step one H2D of the batch, "
                       "LLFQVAE_V4.forward, one D2H of token indices + loss terms; uploads/downloads of neighbouring "
                       "steps overlap the kernels (%d slots)" % DEPTH,
                "host_numa_binding": ("rank bound to %d CPUs local to its GPU" % len(numa_cpus)) if numa_cpus else "none",
                "serial_ms_per_step": ms_e2e_serial,
                "serial_value": world * CHUNKS / (ms_e2e_serial / 1e3)},
        "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "rooflines": rooflines,
        "kernels_us_per_step": kern_us, "cpu_baseline": cpu_baseline,
        "config2_vq_alone": c2, "hbm_kernel": hbm, "train_step": train,
        "parity": parity, "parity_checked_rows": parity["parity_checked_rows"] if parity else 0,
        "latency_us": latency, "config4_sweep": c4, "config5_icrt": c5, "vqvae_classic": vqc,
    }
    _emit(out)


# ------------------------------------------------------------------------------------------ extra legs
def _guard(fn, *a, **k):
    """An extra leg must never cost the headline line: report its error instead."""
    try:
        return fn(*a, **k)
    except Exception as e:           # noqa: BLE001
        return {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}


def spread_params(seed=11):
    """Weights with index diversity (random init collapses every row onto one code): a milder Lipschitz
    layer with both ``min`` branches in use and a codebook sampled from the encoder's own outputs."""
    import numpy as np
    from oracle import lipvq_oracle as orc        # input generator + checker only
    rng = np.random.default_rng(seed)
    p = orc.init_params(F_DIM, D_LAT, K_CODES, H_DIM, seed=seed)
    p["to_latent.W"] = (0.08 * p["to_latent.W"]).astype(np.float32)
    p["to_latent.ci"][::3] = 5.0
    z = orc.encode(p, (2.0 * rng.random((K_CODES, F_DIM)) - 1.0).astype(np.float32))
    p["quantizer.codebook"] = (z + 1e-3 * rng.standard_normal(z.shape)).astype(np.float32)
    return p


def parity_leg(torch, m, p, x_dev, slice_rows=16384):
    """Check outputs of the TIMED forward against the oracle before anything is printed: indices
    of the first `slice_rows` rows of the timed batch bit-exact outside the 1e-5 gap window, z_latent the gathered
    codeword, and -- on a forward of the slice alone -- reconstruction and the loss terms within 1e-5.  Run twice:
    the timed (reference-init, collapsed) weights and spread weights."""
    import numpy as np
    from oracle import lipvq_oracle as orc
    from lipvq_vae_b200 import LLFQVAE_V4
    out = {"rule": "idx bit-exact where the float64 top-2 distance gap > 1e-5 relative; z_latent == E[idx]; "
                   "x_recon, loss terms within 1e-5 (oracle.exact_top2 / oracle.decode on the same rows)"}
    total = 0
    for name, params in (("timed_weights", p), ("spread_weights", spread_params())):
        mm = m
        if name != "timed_weights":
            mm = LLFQVAE_V4(F_DIM, D_LAT, num_codes=K_CODES, hidden_dim=H_DIM).to(x_dev.device)
            mm.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in params.items()})
            mm.search_mode = m.search_mode
        with torch.no_grad():
            z_full, _ = mm(x_dev)                                   # the timed call, whole batch
            idx_full = mm.last_indices[:slice_rows].cpu().numpy()
            z_head = z_full[:slice_rows].cpu().numpy()
            xs_dev = x_dev[:slice_rows].contiguous()
            _, loss_s = mm(xs_dev)                                  # the slice alone (its own loss terms)
            idx_s = mm.last_indices.cpu().numpy()
            terms = mm.last_loss_terms.cpu().numpy().astype(np.float64)
            xr = mm.last_reconstruction.cpu().numpy()
        xs = xs_dev.cpu().numpy()
        E = params["quantizer.codebook"]
        z_e = orc.encode(params, xs)
        ref_idx, d1, d2 = orc.exact_top2(z_e, E)
        must = orc.index_parity_mask(d1, d2)
        ok_idx = bool(np.array_equal(idx_full[must], ref_idx[must]) and np.array_equal(idx_s[must], ref_idx[must]))
        ok_z = bool(np.array_equal(z_head, E[idx_full]))
        xr_code = orc.decode(params, E)
        recon = float(np.sum((xr_code[idx_s].astype(np.float64) - xs) ** 2)) / (slice_rows * F_DIM)
        commit = float(np.sum((E[idx_s].astype(np.float64) - z_e) ** 2)) / (slice_rows * D_LAT)
        ref_loss = recon + 0.5 * commit
        err_loss = abs(float(loss_s) - ref_loss) / abs(ref_loss)
        err_xr = float(orc.rel_err(xr, xr_code[idx_s]))
        ok = ok_idx and ok_z and err_loss < 1e-5 and err_xr < 1e-5 and abs(terms[1] - recon) < 1e-5 * recon
        out[name] = {"rows": slice_rows, "rows_outside_gap_window": int(must.sum()), "unique_codes": int(len(np.unique(ref_idx))),
                     "idx_ok": ok_idx, "z_latent_ok": ok_z, "loss_rel_err": err_loss, "x_recon_rel_err": err_xr, "ok": bool(ok)}
        if not ok:
            raise SystemExit("bench.py: parity check FAILED on %s: %r" % (name, out[name]))
        total += slice_rows
    out["parity_checked_rows"] = total
    return out


def latency_leg(torch, N, dev, p):
    """The reference's real call sizes (N = B*T = 80 rows image config, 500 rows low-dim config; v5:89-93): per-call
    latency of the no-grad forward (rollouts, icl.py:836-843) and of forward+backward (training, icl.py:913-970)."""
    import time as _t
    from lipvq_vae_b200 import LLFQVAE_V4
    m = LLFQVAE_V4(F_DIM, D_LAT, num_codes=K_CODES, hidden_dim=H_DIM).to(dev)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in p.items()})
    res = {"note": "gpu_us = CUDA events around 200 back-to-back calls / 200 (includes launch gaps when the host is the "
                   "limit); kernels_us = sum of this library's kernel durations per call (library CUDA events); "
                   "host_us = wall clock per call with a synchronize at the end; graph_us = CUDA-graph replay of the "
                   "same no-grad call"}
    for n in (80, 500):
        x = torch.randn(n, F_DIM, device=dev)

        def fwd(i):
            with torch.no_grad():
                m(x)

        def fwdbwd(i):
            for q in m.parameters():
                q.grad = None
            _, loss = m(x)
            loss.backward()
        r = {}
        for name, fn in (("forward", fwd), ("forward_backward", fwdbwd)):
            for i in range(10):
                fn(i)
            iters = 200
            gpu_ms = _time_events(fn, iters, torch)
            torch.cuda.synchronize()
            t0 = _t.perf_counter()
            for i in range(iters):
                fn(i)
            torch.cuda.synchronize()
            host_us = 1e6 * (_t.perf_counter() - t0) / iters
            rep = N.profile_kernels(fn, 20)
            r[name] = {"gpu_us": 1e3 * gpu_ms, "host_us": host_us,
                       "kernels_us": 1e3 * sum(v[1] for v in rep.values()) / 20,
                       "launches": sum(v[0] for v in rep.values()) / 20, "kernels": sorted(rep.keys())}
        # CUDA-graph replay of the no-grad forward (static input / output buffers)
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for i in range(3):
                fwd(i)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            with torch.no_grad():
                z_static, loss_static = m(x)
        for i in range(10):
            g.replay()
        r["forward"]["graph_us"] = 1e3 * _time_events(lambda i: g.replay(), 200, torch)
        with torch.no_grad():
            z_eager, loss_eager = m(x)
        r["forward"]["graph_matches_eager"] = bool(torch.equal(z_static, z_eager) and torch.equal(loss_static, loss_eager))
        res["n%d" % n] = r
    return res


def vqvae_classic_leg(torch, dev, rows, steps):
    """the reference's classic VQVAE (vq_vae/backbone.py:6-76; ReLU MLPs, squared distances,
    straight-through estimator), same batch as the headline, reference constructor defaults (K=128)."""
    from lipvq_vae_b200 import VQVAE
    torch.manual_seed(0)
    m = VQVAE(F_DIM, D_LAT).to(dev)
    x = torch.randn(rows, F_DIM, device=dev)

    def fwd(i):
        with torch.no_grad():
            m(x)

    def fwdbwd(i):
        for q in m.parameters():
            q.grad = None
        _, loss = m(x)
        loss.backward()
    out = {"workload": "classic VQVAE (backbone.py:6-76), %d chunks, latent_dim=%d, num_embeddings=128" % (rows // CHUNK_LEN, D_LAT)}
    for name, fn in (("forward", fwd), ("forward_backward", fwdbwd)):
        for i in range(3):
            fn(i)
        ms = _time_events(fn, max(3, steps // 4), torch)
        out[name] = {"ms": ms, "chunks_per_s": rows / CHUNK_LEN / (ms / 1e3)}
    return out


def config4_leg(torch, N, lib, dev, rank, world, peaks, steps):
    """BASELINE config 4: large-codebook sweep, per GPU N=2^20 rows, D in {128,256}, K in {4096,16384,65536}, fp16
    tensor-core distances + fp32 rescoring of near-ties; every rank runs it on its own data (seed 0+rank), times are the
    max over ranks; rank 0 checks a row slice against oracle.exact_top2."""
    import ctypes as C
    import numpy as np
    import torch.distributed as dist
    n = 1 << 20
    sustained = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
    cases = []
    g = torch.Generator(device="cpu").manual_seed(rank)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for D in (128, 256):
        z = torch.rand(n, D, generator=g).to(dev)
        for K in (4096, 16384, 65536):
            E = torch.rand(K, D, generator=g).to(dev)
            need = C.c_size_t(0)
            N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
            ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
            idx = torch.empty(n, dtype=torch.int64, device=dev)
            stats = torch.zeros(4, dtype=torch.int32, device=dev)

            def call(i):
                N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), n, D, K, 1, idx.data_ptr(), None,
                                            stats.data_ptr(), ws.data_ptr(), ws.numel(), st))
            for i in range(2):
                call(i)
            ms = _time_events(call, max(3, steps // 6), torch)
            rep = N.profile_kernels(call, 2)
            if world > 1:
                t = torch.tensor([ms], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            tf = n * 2.0 * K * D / (ms / 1e3) / 1e12
            c = {"D": D, "K": K, "rows_per_gpu": n, "call_ms": ms, "call_tflops": tf, "frac_of_sustained_bf16": tf / sustained,
                 "aggregate_tflops": world * tf, "fallback_rows": int(stats[0].item()), "rescored_codes": int(stats[1].item()),
                 "kernels_us": {k: 1e3 * v[1] / 2 for k, v in rep.items()}}
            if "vq_tc_kernel" in rep:
                c["vq_tc_kernel_tflops"] = n * 2.0 * K * D / (rep["vq_tc_kernel"][1] / 2 / 1e3) / 1e12
            if rank == 0:
                from oracle import lipvq_oracle as orc
                rows = 16384 if K <= 16384 else 4096
                ref, d1, d2 = orc.exact_top2(z[:rows].cpu().numpy(), E.cpu().numpy(), block_rows=512 if K > 16384 else 2048)
                must = orc.index_parity_mask(d1, d2)
                got = idx[:rows].cpu().numpy()
                c["parity_rows"] = rows
                c["parity_ok"] = bool(np.array_equal(got[must], ref[must]))
                if not c["parity_ok"]:
                    raise SystemExit("bench.py: config 4 parity FAILED at D=%d K=%d" % (D, K))
            cases.append(c)
            del E, ws, idx
        del z
    return {"workload": "large-codebook sweep, N=2^20 rows per GPU, fp16 tensor-core distances + fp32 near-tie rescore "
                        "(lipvq_vq_argmin from fp32 z / E, operand conversion and resolve passes included)",
            "n_gpus": world, "peak": sustained, "peak_source": "bf16 sustained, MEASURED_PEAKS.json", "cases": cases,
            "timing": "CUDA events, max over ranks"}


def config5_leg(torch, dev, rank, world, steps):
    """BASELINE config 5: ICRT in-context transformer policy training with the LipVQ tokenizer inline (icl.py:891-974
    replayed by lipvq_vae_b200.icrt.ICRTTrainer; the policy is stock torch, pinned to the reference network by
    tests/test_icrt_integration_cpu.py), synthetic prompt/query sequences, data parallel (weak scaling)."""
    from collections import OrderedDict
    import torch.distributed as dist
    from lipvq_vae_b200.icrt import ICRTPolicy, ICRTTrainer, synthetic_icrt_batch
    obs_shapes = OrderedDict(robot0_eef_pos=[3], robot0_eef_quat=[4], robot0_gripper_qpos=[2], object=[199])   # D = 208
    torch.manual_seed(0)
    pol = ICRTPolicy(obs_shapes, embed_dim=512, num_layers=6, num_heads=8, context_length=10).to(dev).train()
    if world > 1:
        for q in pol.parameters():
            dist.broadcast(q.data, src=0)
    tr = ICRTTrainer(pol)
    out = {"workload": "ICRT training iteration (tokenizer fwd+bwd+AdamW, then policy NLL fwd+bwd+AdamW), %.1f M-parameter "
                       "policy (6 layers, d=512, 30 tokens), tokenizer D=208 K=1024, synthetic sequences; "
                       "policy = stock torch eager fp32" % (sum(q.numel() for q in pol.parameters()) / 1e6),
           "n_gpus": world, "cases": []}
    for n_seq in (32, 512, 4096):          # B = n_seq/2 query sequences (+ as many prompts): 160 / 2560 / 20480 tokenizer rows
        batch = synthetic_icrt_batch(obs_shapes, n_seq, device=dev, seed=100 + rank)
        for i in range(2):
            tr.step(batch)
        k = max(3, steps // 4)
        ms = _time_events(lambda i: tr.step(batch), k, torch)
        # the tokenizer's share: the same rows through forward + backward + its AdamW alone
        tok, opt = pol.vq_vae_model, tr.vq_optimizer
        rows = batch["actions"][:n_seq // 2].reshape(-1, 12).contiguous()

        def tok_only(i):
            opt.zero_grad()
            _, l = tok(rows)
            l.backward()
            opt.step()
        for i in range(2):
            tok_only(i)
        ms_tok = _time_events(tok_only, k, torch)
        if world > 1:
            t = torch.tensor([ms, ms_tok], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, ms_tok = t.tolist()
        out["cases"].append({"sequences_per_gpu": n_seq, "tokenizer_rows_per_gpu": int(rows.shape[0]), "ms_per_iteration": ms,
                             "iterations_per_s": 1e3 / ms, "sequences_per_s": world * n_seq / (ms / 1e3),
                             "tokenizer_ms": ms_tok, "tokenizer_share": ms_tok / ms})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--search", default="tensor", choices=["tensor", "exact"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--skip", default="", help="comma list of extra legs to skip: parity,latency,config4,config5")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # libraries (NCCL's version banner, ...) write to fd 1: keep stdout for the ONE JSON line
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
