"""Prompt-token cache for rollouts (host side).

During evaluation rollouts the reference re-tokenises the SAME in-context prompt on every environment step:
``rollout_with_stats`` (robomimic/utils/train_utils.py, called from scripts/train.py:354-364) calls the policy once
per step, and ``ICLObservationGroupEncoder.forward`` runs ``self.action_network(prompt_actions)`` each time
(obs_nets.py:1336) although neither the prompt nor the tokenizer parameters change inside an episode
(algo/icl.py:827-853 runs under ``torch.no_grad()``).  The tokenizer is deterministic, so its output for an unchanged
prompt and unchanged parameters can be replayed.

``PromptTokenCache`` wraps the tokenizer module and keeps ``(z_latent, loss, indices)`` per prompt:

* key of a HOST batch: BLAKE2b of its bytes (a prompt is 80-500 rows x 12 floats: microseconds to hash, and the
  upload is skipped on a hit as well);
* key of a DEVICE batch: ``(data_ptr, shape, stride, dtype, _version)`` with a reference to the tensor held so its
  memory cannot be recycled under the key; an in-place write bumps ``_version`` and misses;
* any in-place parameter update (optimizer step, ``load_state_dict``) bumps a parameter ``_version`` and clears the
  cache;
* with autograd enabled the call goes straight to the module (the training step needs the graph, and its prompts
  differ every step anyway) -- the wrapper never changes what training computes.

The wrapper is host logic only: every miss runs the module's own CUDA path; there is no CPU tokenizer here.
"""
import collections
import hashlib

import torch

__all__ = ["PromptTokenCache"]


class PromptTokenCache(torch.nn.Module):
    def __init__(self, model: torch.nn.Module, capacity: int = 64):
        super().__init__()
        if capacity < 1:
            raise ValueError("capacity must be >= 1")
        self.model = model
        self.capacity = int(capacity)
        self._entries = collections.OrderedDict()     # key -> (z_latent, loss, indices, pinned input or None)
        self._param_state = None
        self.hits = 0
        self.misses = 0

    # the wrapped module's parameters / state_dict stay reachable under the reference's names
    # (``vq_vae_model.parameters()`` at icl.py:886-889 sees the same 14 tensors through ``self.model``)
    def _versions(self):
        return tuple((p.data_ptr(), p._version) for p in self.model.parameters())

    @staticmethod
    def _key(x: torch.Tensor):
        if x.device.type == "cpu":
            xc = x.detach().contiguous()
            h = hashlib.blake2b(xc.numpy().tobytes(), digest_size=16).digest()
            return ("host", tuple(xc.shape), str(xc.dtype), h)
        return ("dev", x.device.index, x.data_ptr(), tuple(x.shape), tuple(x.stride()), str(x.dtype), x._version)

    def clear(self):
        self._entries.clear()

    def forward(self, x: torch.Tensor):
        if torch.is_grad_enabled():
            return self.model(x)
        state = self._versions()
        if state != self._param_state:
            self._entries.clear()
            self._param_state = state
        key = self._key(x)
        hit = self._entries.get(key)
        if hit is not None:
            self._entries.move_to_end(key)
            self.hits += 1
            self.last_indices = hit[2]
            return hit[0], hit[1]
        self.misses += 1
        xin = x
        if x.device.type == "cpu":
            dev = next(self.model.parameters()).device
            xin = x.to(dev, non_blocking=True) if dev.type != "cpu" else x
        z_latent, loss = self.model(xin)
        idx = getattr(self.model, "last_indices", None)
        self.last_indices = idx
        # a device key is only valid while the tensor's storage is alive: keep it
        self._entries[key] = (z_latent, loss, idx, x if x.device.type != "cpu" else None)
        while len(self._entries) > self.capacity:
            self._entries.popitem(last=False)
        return z_latent, loss
