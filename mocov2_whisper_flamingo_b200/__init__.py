"""B200-native (sm_100a) gated cross-attention path of MoCov2-Whisper-Flamingo.

Public surface == the reference's models/gate_cross_attention.py: ``LayerNorm``,
``GatedCrossAttentionBlock``, ``GatedCrossModalFusion`` (same ctor / forward / state_dict).
Everything below them is hand-written CUDA behind the C ABI in include/gxa.h.
"""
from .gate_cross_attention import (GatedCrossAttentionBlock, GatedCrossModalFusion, LayerNorm,  # noqa: F401
                                   set_default_precision)
from .decode import DecodeState  # noqa: F401
from .av_glue import fuse_sum, positional_table, proj_ln_pe  # noqa: F401
from ._lib import LIB_PATH, launch_count  # noqa: F401

__all__ = ["LayerNorm", "GatedCrossAttentionBlock", "GatedCrossModalFusion", "set_default_precision",
           "DecodeState", "proj_ln_pe", "fuse_sum", "positional_table", "launch_count", "LIB_PATH"]
