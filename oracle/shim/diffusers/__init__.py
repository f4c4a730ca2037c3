"""Minimal stand-in for the `diffusers` package (TEST INFRASTRUCTURE ONLY).

The reference (HuaizeLiu/Hi-VAE) imports diffusers 0.30.0 (version recorded at
/root/reference/audio2motion/config.json:3), which is not installed and cannot be
installed offline.  This shim restates, in plain PyTorch, only the names the
reference's hot path imports (model/modules.py:24-35, model/transformer.py:7-32,
model/model_AMD.py:23-30) so that the reference's own files import UNMODIFIED and
run in fp32 on CPU.  Parity status: the reference ships no golden vectors, so the
restated semantics are "parity unpinned" against real diffusers.
"""


class CogVideoXPipeline:  # imported, never used on the hot path
    pass


class CogVideoXDDIMScheduler:
    pass


class AutoencoderKL:
    pass
