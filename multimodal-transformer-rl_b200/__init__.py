"""multimodal-transformer-rl_b200 — B200-native (sm_100a) Decision-Transformer hot path.

Python host mirror of the reference API (``DecisionTransformer``, ``SequenceTrainer``,
``MultimodalStateEncoder``) over the C ABI in ``include/dtb200.h`` (``libdtb200.so``: hand-written
tcgen05 / TMA / TMEM CUDA kernels).  There is no CPU or eager-PyTorch fallback: constructing a model
without the built library or without an sm_100 device raises.
Import as ``mtrl_b200`` (the directory name is not a valid Python identifier).
"""
__version__ = "0.1.0"

__all__ = ["__version__"]


def __getattr__(name):  # lazy: importing the package must not require torch.cuda
    import importlib

    lazy = {
        "DecisionTransformer": "model",
        "SequenceTrainer": "trainer",
        "MultimodalStateEncoder": "encoder",
        "MultimodalDecisionTransformer": "encoder",
        "lib": "_lib",
    }
    if name in lazy:
        mod = importlib.import_module("mtrl_b200." + lazy[name])
        return getattr(mod, name) if name != "lib" else mod
    raise AttributeError(name)
