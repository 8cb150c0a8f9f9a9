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
    _emit({
        "impl": "reference", "metric": "action chunks tokenized/sec (encode+quantize+decode)", "value": val,
        "unit": "chunks/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": val, "unit": "chunks/s", "cores": cores, "kind": "port", "sample": sample},
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
    tr_steps = max(3, args.steps // 4)

    def train_step(i):
        trainer.step(xs[i % NBUF], n_total)
    for i in range(2):
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
    rep = N.profile_kernels(train_step, 3)
    torch.cuda.synchronize()
    barrier()
    if rank == 0:
        # the two backward GEMM families report against their rooflines: weight gradients 2*N*sum(Nout*Kin) flop as
        # 3 tf32 MMAs per product, input gradients likewise; both also move every saved activation once
        train["kernels_us_per_step"] = {k: 1e3 * v[1] / 3 for k, v in rep.items()}
        Hh, H1h, D, F = H_DIM, 64, D_LAT, F_DIM
        hbm_peak_for_train = _peaks()[0]["hbm_gbs"]
        wg_flop = 2.0 * rows * (D * Hh + Hh * H1h + H1h * F)
        dg_flop = 2.0 * rows * (D * Hh + Hh * H1h)
        wg_bytes = 4.0 * rows * (D + Hh + Hh + H1h + H1h + F)
        dg_bytes = 4.0 * rows * (D + 2 * Hh + Hh + 2 * H1h)
        for kname, flop, byts in (("wgrad_umma_kernel", wg_flop, wg_bytes), ("dgrad_umma_kernel", dg_flop, dg_bytes)):
            us = train["kernels_us_per_step"].get(kname)
            if us:
                train.setdefault("backward_rooflines", []).append({
                    "kernel": kname, "us_per_step": us, "bound": "hbm", "achieved": byts / (us * 1e-6) / 1e9,
                    "peak": hbm_peak_for_train, "unit": "GB/s", "frac": byts / (us * 1e-6) / 1e9 / hbm_peak_for_train,
                    "tensor_tflops_fp32_equivalent": flop / (us * 1e-6) / 1e12,
                    "algorithmic": "reads of dY and the saved activations + writes of the activation gradients, fp32, once each"})
    del trainer, m_tr
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

    if rank != 0:
        return
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
        _roof("enc_fused_kernel", "hbm", rows * (F_DIM * 4.0 + D_LAT * 4 + ZB_BYTES), peaks["hbm_gbs"], "GB/s",
              "read x (F*4 B) + write z_e (D*4 B) + write the search's fp16 row operand (%d B) per row; its GEMMs are "
              "71168 fp32-equivalent flop/row run as 3 kind::f16 MMAs per product (tensor floor 98 us < HBM floor "
              "133 us), see `tensor_tflops_fp32_equivalent`" % ZB_BYTES),
        _roof("quantize_emit_kernel", "hbm", rows * (2.0 * D_LAT * 4 + 8 + 2 * F_DIM * 4), peaks["hbm_gbs"], "GB/s",
              "read z_e + write z_q (2*D*4 B/row) + idx 8 B + x read and x_recon write (2*F*4 B)"),
    ) if r]
    # dram bytes per launch of the same kernels on the same workload, from the committed ncu --set full capture
    try:
        _traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))["kernels"]
    except Exception:
        _traffic = {}
    for r in rooflines:
        t_ = _traffic.get(r["kernel"])
        r["traffic"] = t_["total_bytes"] if t_ else None
        r["traffic_source"] = "profiles/r01_traffic.json (ncu dram__bytes_read.sum + dram__bytes_write.sum)" if t_ else None
    for r in rooflines:
        r["peak_source"] = peak_src + (" bf16 sustained (kernel timed inside the step)" if r["bound"] == "tensor"
                                       else " HBM copy bandwidth")
    for r in rooflines:
        if r["kernel"] == "enc_fused_kernel":
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
                "api": "HostTokenizer.submit(pinned host batch) / collect(): per step one H2D of the batch, "
                       "LLFQVAE_V4.forward, one D2H of token indices + loss terms; uploads/downloads of neighbouring "
                       "steps overlap the kernels (%d slots)" % DEPTH,
                "host_numa_binding": ("rank bound to %d CPUs local to its GPU" % len(numa_cpus)) if numa_cpus else "none",
                "serial_ms_per_step": ms_e2e_serial,
                "serial_value": world * CHUNKS / (ms_e2e_serial / 1e3)},
        "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "rooflines": rooflines,
        "kernels_us_per_step": kern_us, "cpu_baseline": cpu_baseline,
        "config2_vq_alone": c2, "hbm_kernel": hbm, "train_step": train,
    }
    _emit(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--search", default="tensor", choices=["tensor", "exact"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
