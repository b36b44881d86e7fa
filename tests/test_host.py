"""CPU-side checks: the C-ABI library loads and exports every symbol include/gxa.h declares, the host
shell mirrors the reference's module interface (names, ctor args, state_dict layout, error types) and
refuses to run without CUDA (no CPU fallback)."""
import ctypes as C
import os
import re
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import gxa_oracle as O
import mocov2_whisper_flamingo_b200 as G
from mocov2_whisper_flamingo_b200 import _lib


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gxa.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gxa_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build with python -m mocov2_whisper_flamingo_b200.build"
    L = C.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/gxa.h but not exported"
    assert set(syms) == set(_lib.SIGNATURES), "ctypes signatures out of sync with include/gxa.h"
    assert b"sm_100a" in _lib.lib().gxa_version()


def test_size_queries_are_host_only_and_consistent():
    L = _lib.lib()
    d = _lib.Dims(B=2, Lq=448, Lk=250, D=384, H=6, precision=0, has_xa=1, reserved=0)
    assert L.gxa_block_wcache_bytes(C.byref(d)) == 12 * 384 * 384 * 2
    sv, ws = L.gxa_block_saved_bytes(C.byref(d)), L.gxa_block_workspace_bytes(C.byref(d))
    assert sv > 2 * 448 * 384 * (4 + 2 * 5 + 2 * 8) and ws > 0
    bad = _lib.Dims(B=2, Lq=448, Lk=250, D=384, H=5, precision=0, has_xa=1, reserved=0)   # head_dim != 64
    assert L.gxa_block_saved_bytes(C.byref(bad)) == 0
    assert L.gxa_block_fwd(C.byref(bad), None, None, None, None, None, None, None, None, None, None) == -1


def test_state_dict_layout_and_strict_load():
    blk = G.GatedCrossAttentionBlock(128, 2, dropout=0.0)
    assert list(blk.state_dict().keys()) == list(O.BLOCK_KEYS)
    assert {k: tuple(v.shape) for k, v in blk.state_dict().items()} == O.block_param_shapes(128)
    fus = G.GatedCrossModalFusion(128, 2, 3)
    shapes = O.fusion_param_shapes(128, 3)
    assert list(fus.state_dict().keys()) == list(shapes.keys())
    p = O.make_params(shapes, 3)
    fus.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    assert fus.layers[1].attn_gate.item() == pytest.approx(0.7)        # gates stay readable Parameters (monitor.py)
    assert fus.layers[1].ff_gate.item() == pytest.approx(-0.4)


def test_init_matches_reference_defaults():
    blk = G.GatedCrossAttentionBlock(128, 2)
    assert blk.attn_gate.item() == 0.0 and blk.ff_gate.item() == 0.0     # identity at init (A.1 M6)
    assert float(blk.attn.in_proj_bias.detach().abs().max()) == 0.0 and float(blk.attn.out_proj.bias.detach().abs().max()) == 0.0
    assert tuple(blk.ff[0].weight.shape) == (512, 128)                   # hidden = 4 * d_model


def test_error_conventions():
    with pytest.raises(AssertionError, match="embed_dim must be divisible by num_heads"):
        G.GatedCrossAttentionBlock(100, 3)
    blk = G.GatedCrossAttentionBlock(128, 2, dropout=0.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        blk(torch.randn(1, 4, 128), torch.randn(1, 4, 128))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        G.LayerNorm(128)(torch.randn(2, 128))
    with pytest.raises(RuntimeError, match="no CPU fallback"):       # training-mode dropout is on the CUDA path too
        G.GatedCrossAttentionBlock(128, 2, dropout=0.1).train()(torch.randn(1, 4, 128))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "mocov2_whisper_flamingo_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, f)).read(), f


def test_dropout_mask_restatement_statistics():
    """tests/dropout_ref.py (numpy restatement of csrc/gxa_rng.cuh): drop rate, independence, seed sensitivity."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    import numpy as np
    import dropout_ref as DR
    m = DR.multipliers(512, 751, 0.1, 0x1234ABCD5678EF01, DR.STREAM_ATTN)
    keep = m > 0
    assert set(np.unique(m)) == {np.float32(0.0), np.float32(1.0) / (np.float32(1.0) - np.float32(0.1))}
    assert abs(keep.mean() - 0.9) < 2e-3
    assert abs(np.corrcoef(keep[:, :-1].ravel(), keep[:, 1:].ravel())[0, 1]) < 5e-3      # neighbouring keys
    assert abs(np.corrcoef(keep[:-1].ravel(), keep[1:].ravel())[0, 1]) < 5e-3            # neighbouring rows
    m2 = DR.multipliers(512, 751, 0.1, 0x1234ABCD5678EF02, DR.STREAM_ATTN)
    assert abs(np.corrcoef(keep.ravel(), (m2 > 0).ravel())[0, 1]) < 5e-3                 # another seed
    m3 = DR.multipliers(512, 751, 0.1, 0x1234ABCD5678EF01, DR.STREAM_FFN)
    assert abs(np.corrcoef(keep.ravel(), (m3 > 0).ravel())[0, 1]) < 5e-3                 # another stream
    assert np.all(DR.multipliers(4, 9, 0.0, 1, 1) == 1)


def test_balanced_shards_even_out_frame_counts():
    """length-sorted serpentine assignment vs contiguous shards at config-4 lengths (60..750 frames)."""
    import random
    from mocov2_whisper_flamingo_b200.parallel import balanced_shards, shard_batch
    rng = random.Random(1234)
    lengths = sorted(rng.randint(60, 750) for _ in range(64))          # worst case for contiguous shards: sorted input
    for world in (1, 2, 4, 8):
        shards = balanced_shards(lengths, world)
        assert sorted(i for s in shards for i in s) == list(range(64))                 # a partition
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
        tot = [sum(lengths[i] for i in s) for s in shards]
        contiguous = [sum(lengths[slice(*shard_batch(64, r, world))]) for r in range(world)]
        assert max(tot) - min(tot) <= 750                                              # within one sample's length
        assert max(tot) <= max(contiguous)
        if world > 1:
            assert max(tot) / (sum(tot) / world) < 1.05 < 1.3 < max(contiguous) / (sum(contiguous) / world)
    assert balanced_shards([5, 5, 5], 2) == [[0], [1, 2]]                              # serpentine; ties keep index order
    assert balanced_shards([], 3) == [[], [], []]


def test_dropout_restatement_is_bit_identical_to_the_compiled_mask_function():
    """tests/dropout_ref.py (what the -m gpu parity tests feed the oracle) against the mask function the kernels are
    compiled from (csrc/gxa_rng.cuh), evaluated on the host through gxa_dropout_multipliers_host: every multiplier equal."""
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    import numpy as np
    import dropout_ref as DR
    L = _lib.lib()
    for rows, cols, p, seed, site in [(96, 750, 0.1, 0x1234ABCD5678EF01, DR.STREAM_ATTN),
                                      (37, 751, 0.1, 0xFFFFFFFFFFFFFFFF, DR.STREAM_ATTN),      # odd column count
                                      (64, 1024, 0.25, 7, DR.STREAM_FFN),
                                      (5, 9, 0.999, 3, DR.STREAM_FFN),
                                      (3, 8, 0.0, 11, DR.STREAM_ATTN)]:
        out = np.empty((rows, cols), np.float32)
        rc = L.gxa_dropout_multipliers_host(p, seed, site, rows, cols, out.ctypes.data)
        assert rc == 0
        ref = DR.multipliers(rows, cols, p, seed, site)
        assert np.array_equal(out, ref), (rows, cols, p, site)
    assert L.gxa_dropout_multipliers_host(1.0, 1, 1, 1, 1, np.empty(1, np.float32).ctypes.data) != 0      # p must be < 1
    assert L.gxa_dropout_multipliers_host(0.1, 1, 3, 1, 1, np.empty(1, np.float32).ctypes.data) != 0      # unknown site


def test_bench_reference_arm_prints_exactly_one_json_line():
    """bench.py's contract: ONE JSON line on stdout, whatever native libraries print (descriptor 1 is re-pointed at stderr
    for the run).  The reference arm runs on the host cores, so this part of the contract is checkable without a GPU; a
    child process that writes to descriptor 1 behind Python's back stands in for NCCL's banner."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import os, sys, bench; sys.argv = ['bench.py', '--impl', 'reference', '--workload', 'whisper-tiny-c1', "
            "'--steps', '1', '--warmup', '0', '--ref-batch', '1']; "
            "bench.claim_stdout(); os.write(1, b'NCCL version 0.0.0 (stand-in)\\n'); "
            "bench.claim_stdout = lambda: None; bench.main()")
    r = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    out = json.loads(lines[0])
    assert out["impl"] == "reference" and out["cpu_baseline"]["kind"] in ("reference", "port")
    assert out["e2e"]["h2d_bytes_per_step"] == 0 and out["value"] > 0
    assert "NCCL version" in r.stderr
