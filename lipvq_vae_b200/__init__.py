"""lipvq_vae_b200 -- B200-native LipVQ-VAE action tokenizer (one hot path of andvg3/LipVQ-VAE).

Public surface mirrors the reference module robomimic/models/vq_vae/backbone_lfqvae_v5.py:
``LLFQVAE_V4`` (+ ``LipschitzMLP``, ``LFQQuantizer``) backed by hand-written sm_100a kernels
behind the C ABI in include/lipvq_b200.h.  No CPU path, no torch fallback.
"""
from ._native import (LipVQNativeError, SEARCH_EXACT, SEARCH_TENSOR, build_native, lib as native_lib,
                      LIB_PATH)
from .tokenizer import LLFQVAE_V4, LFQQuantizer, LipschitzMLP, install_into_robomimic
from .vqvae_classic import VQVAE
from .stream import HostTokenizer, bind_host_thread_to_gpu_node
from .dist import DataParallelTokenizer, FlatAdamW, TokenizerTrainer, shard_rows
from .cache import PromptTokenCache

__all__ = ["LLFQVAE_V4", "VQVAE", "LFQQuantizer", "LipschitzMLP", "install_into_robomimic", "HostTokenizer", "bind_host_thread_to_gpu_node", "DataParallelTokenizer", "FlatAdamW", "TokenizerTrainer",
           "shard_rows", "PromptTokenCache", "LipVQNativeError",
           "SEARCH_EXACT", "SEARCH_TENSOR", "build_native", "native_lib", "LIB_PATH"]
