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
  cache; updates that write parameters through raw pointers (``FlatAdamW.step``, ``ema_update``,
  ``TokenizerTrainer.step``) advance the model's ``_param_epoch`` instead, which is part of the same check;
* with autograd enabled the call goes straight to the module (the training step needs the graph, and its prompts
  differ every step anyway) -- the wrapper never changes what training computes.

The wrapper is host logic only: every miss runs the module's own CUDA path; there is no CPU tokenizer here.
"""
import collections
import hashlib

import torch

__all__ = ["PromptTokenCache"]


class PromptTokenCache(torch.nn.Module):
    """Wraps the module that is actually CALLED on the prompt -- ``nets["policy"].nets["encoder"].action_network``
    (robomimic/models/obs_nets.py:1336); ``vq_vae_model`` (obs_nets.py:2422, 2721) is only an alias of it.

    The wrapper is state-dict transparent: the inner module's children (``encoder``, ``to_latent``, ``quantizer``,
    ``decoder``, ``to_output``) and direct parameters are registered on the wrapper under the SAME names, the inner
    module itself is held outside ``_modules``.  ``state_dict()`` / strict ``load_state_dict()`` /
    ``parameters()`` / ``.to()`` therefore see exactly the reference's 14 keys (robomimic/algo/algo.py:323-337), with
    no ``model.`` prefix."""

    def __init__(self, model: torch.nn.Module, capacity: int = 64):
        super().__init__()
        if capacity < 1:
            raise ValueError("capacity must be >= 1")
        object.__setattr__(self, "_inner", model)          # not a registered child: no key prefix
        # (written into the registries directly: add_module()'s hasattr check would find the inner module's attribute
        #  of the same name through __getattr__ below)
        for name, child in model.named_children():
            self._modules[name] = child                    # same objects, same names as the reference class
        for name, prm in model.named_parameters(recurse=False):
            self._parameters[name] = prm
        for name, buf in model.named_buffers(recurse=False):
            self._buffers[name] = buf
        self.capacity = int(capacity)
        self._entries = collections.OrderedDict()     # key -> (z_latent, loss, indices, pinned input or None)
        self._param_state = None
        self.hits = 0
        self.misses = 0

    @property
    def model(self):
        return self._inner

    def __getattr__(self, name):
        # attributes of the wrapped tokenizer (last_indices, search_mode, encode/quantize/decode ...) stay reachable
        try:
            return super().__getattr__(name)
        except AttributeError:
            inner = self.__dict__.get("_inner")
            if inner is None:
                raise
            return getattr(inner, name)

    def _versions(self):
        """Identity + autograd version of every parameter, plus the model's parameter epoch: updates that go through
        raw pointers (FlatAdamW.step, ema_update, TokenizerTrainer.step, the data-parallel broadcast) never bump
        ``_version`` and advance ``_param_epoch`` instead."""
        m = self._inner
        return (getattr(m, "_param_epoch", 0),) + tuple((p.data_ptr(), p._version) for p in m.parameters())

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
        m = self._inner
        if torch.is_grad_enabled():
            return m(x)
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
            dev = next(m.parameters()).device
            xin = x.to(dev, non_blocking=True) if dev.type != "cpu" else x
        z_latent, loss = m(xin)
        idx = getattr(m, "last_indices", None)
        self.last_indices = idx
        # a device key is only valid while the tensor's storage is alive: keep it
        self._entries[key] = (z_latent, loss, idx, x if x.device.type != "cpu" else None)
        while len(self._entries) > self.capacity:
            self._entries.popitem(last=False)
        return z_latent, loss
