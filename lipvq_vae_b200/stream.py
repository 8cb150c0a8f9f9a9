"""Host-batch tokenisation pipeline: the serving-side entry point for batches that live in pinned HOST
memory (the ICRT rollout / data-loader side hands over CPU tensors, robomimic/utils/train_utils.py).

``HostTokenizer.submit(x_host)`` enqueues, without blocking the host,
    H2D copy of the batch (copy-in stream)  ->  LLFQVAE_V4.forward (the caller's current stream)
    ->  D2H copy of the token indices and the four loss terms (copy-out stream)
into one of ``depth`` slots; ``collect()`` returns the oldest outstanding batch's results from pinned host
buffers.  With ``depth >= 2`` the upload of batch i+1 and the download of batch i-1 overlap the kernels of
batch i (B200: the copy engines run beside the SMs), so the steady-state step time is
max(copy, compute) instead of their sum.  z_latent stays on the device (it feeds the policy there,
robomimic/models/obs_nets.py:1336); pass ``fetch_latent=True`` to bring it back as well.
"""
from __future__ import annotations

import collections

import torch

from .tokenizer import LLFQVAE_V4, LipVQNativeError


def bind_host_thread_to_gpu_node(device_index: int):
    """Pin the calling process to the CPUs local to ``cuda:device_index`` (sysfs ``local_cpulist`` of the GPU's PCI
    function) so that pinned host buffers allocated afterwards sit on the GPU's own NUMA node: with one process per
    GPU, batches staged on the wrong socket cross the inter-socket link on every upload.  Returns the CPU list, or
    ``None`` when the topology is not exposed (single node, containers without sysfs) -- nothing is changed then."""
    import os
    try:
        pr = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


class HostTokenizer:
    def __init__(self, model: LLFQVAE_V4, max_rows: int, depth: int = 2, fetch_latent: bool = False):
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise LipVQNativeError("HostTokenizer needs the model on a CUDA device (there is no CPU path)")
        self.model, self.dev, self.depth, self.max_rows = model, dev, int(depth), int(max_rows)
        self.fetch_latent = fetch_latent
        F, D = model.feature_dim, model.latent_dim
        self.copy_in = torch.cuda.Stream(device=dev)
        self.copy_out = torch.cuda.Stream(device=dev)
        self.x_dev = [torch.empty(max_rows, F, device=dev) for _ in range(depth)]
        self.idx_host = [torch.empty(max_rows, dtype=torch.int64).pin_memory() for _ in range(depth)]
        self.loss_host = [torch.empty(4).pin_memory() for _ in range(depth)]
        self.lat_host = [torch.empty(max_rows, D).pin_memory() for _ in range(depth)] if fetch_latent else None
        self.ev_in = [torch.cuda.Event() for _ in range(depth)]
        self.ev_done = [torch.cuda.Event() for _ in range(depth)]      # forward finished: x_dev[slot] reusable
        self.ev_out = [torch.cuda.Event() for _ in range(depth)]
        self.pending = collections.deque()                             # (slot, rows, z_latent)
        self.n_submitted = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def submit(self, x_host: torch.Tensor):
        """Enqueue one pinned-host batch [rows, feature_dim]; returns its z_latent (device tensor)."""
        if x_host.is_cuda:
            raise ValueError("submit() takes a host tensor; call the model directly for device batches")
        rows = x_host.shape[0]
        if rows > self.max_rows:
            raise ValueError("batch of %d rows exceeds max_rows=%d" % (rows, self.max_rows))
        if len(self.pending) >= self.depth:
            raise RuntimeError("all %d slots are in flight: collect() first" % self.depth)
        slot = self.n_submitted % self.depth
        self.n_submitted += 1
        main = torch.cuda.current_stream(self.dev)
        xd = self.x_dev[slot][:rows]
        self.copy_in.wait_event(self.ev_done[slot])          # the slot's previous forward has consumed x_dev
        with torch.cuda.stream(self.copy_in):
            xd.copy_(x_host, non_blocking=True)
            self.ev_in[slot].record(self.copy_in)
        main.wait_event(self.ev_in[slot])
        with torch.no_grad():
            z_latent, _ = self.model(xd)
        idx, terms = self.model.last_indices, self.model.last_loss_terms
        self.ev_done[slot].record(main)
        self.copy_out.wait_event(self.ev_done[slot])
        with torch.cuda.stream(self.copy_out):
            self.idx_host[slot][:rows].copy_(idx, non_blocking=True)
            self.loss_host[slot].copy_(terms, non_blocking=True)
            if self.fetch_latent:
                self.lat_host[slot][:rows].copy_(z_latent, non_blocking=True)
                z_latent.record_stream(self.copy_out)
            idx.record_stream(self.copy_out)
            terms.record_stream(self.copy_out)
            self.ev_out[slot].record(self.copy_out)
        self.h2d_bytes += x_host.numel() * 4
        self.d2h_bytes += rows * 8 + 16 + (z_latent.numel() * 4 if self.fetch_latent else 0)
        self.pending.append((slot, rows, z_latent))
        return z_latent

    def collect(self):
        """Block until the oldest outstanding batch is back on the host; returns
        ``(indices int64[rows], loss_terms float[4] = loss/recon/commit/codebook, z_latent)`` -- the first two
        are views of pinned host buffers that the slot's next ``submit`` overwrites."""
        if not self.pending:
            raise RuntimeError("nothing in flight")
        slot, rows, z_latent = self.pending.popleft()
        self.ev_out[slot].synchronize()
        lat = self.lat_host[slot][:rows] if self.fetch_latent else z_latent
        return self.idx_host[slot][:rows], self.loss_host[slot], lat

    def drain_to(self, stream=None):
        """Make ``stream`` (default: the current one) wait for every enqueued download (for device-side timing)."""
        stream = stream or torch.cuda.current_stream(self.dev)
        for slot, _, _ in self.pending:
            stream.wait_event(self.ev_out[slot])
