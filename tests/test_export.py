"""SURVEY 8 f4 -- the ONNX graph the export emits (mocov2_whisper_flamingo_b200/export.py), checked on the CPU.

``torch.onnx`` traces by running the module, and the product's forward needs a GPU.  What is under test here is the graph
EMISSION (the ``symbolic`` functions), so at trace time the three export Functions' ``forward`` are stubbed with the
oracle's torch port (the checker; the tracer only needs tensors of the right shape from them).  The emitted graph is then
evaluated node by node (tests/onnx_interp.py, float64) and compared with the golden outputs of the live reference module
and, at shapes other than the traced ones (dynamic axes), with the torch port.  tests/test_export_gpu.py repeats the
comparison with the real CUDA forward.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import gxa_oracle as O
from oracle import torch_port as TP
import mocov2_whisper_flamingo_b200 as G
from mocov2_whisper_flamingo_b200 import export as E
import onnx_interp

KEYS = ("attn_gate", "ff_gate", "attn.in_proj_weight", "attn.in_proj_bias", "attn.out_proj.weight", "attn.out_proj.bias",
        "attn_ln.weight", "attn_ln.bias", "ff_ln.weight", "ff_ln.bias", "ff.0.weight", "ff.0.bias", "ff.2.weight",
        "ff.2.bias")


@pytest.fixture
def stub_forward(monkeypatch):
    def block_fwd(ctx, x, xa, mask, n_heads, block, *params):
        return TP.block_forward(dict(zip(KEYS, params)), x, xa, mask, n_heads)

    def linear_fwd(ctx, x, w, b, precision):
        return torch.nn.functional.linear(x, w, b)

    def ln_fwd(ctx, x, w, b):
        return torch.nn.functional.layer_norm(x, x.shape[-1:], w, b)

    monkeypatch.setattr(E._BlockExportFn, "forward", staticmethod(block_fwd))
    monkeypatch.setattr(E._LinearExportFn, "forward", staticmethod(linear_fwd))
    monkeypatch.setattr(E._LayerNormExportFn, "forward", staticmethod(ln_fwd))


def _load(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    B, Lq, Lk, D, H, nl, seed = (int(v) for v in g["spec"])
    kind = "fusion" if name.startswith("fusion") else ("noxa" if "noxa" in name else "block")
    shapes = O.fusion_param_shapes(D, nl) if kind == "fusion" else O.block_param_shapes(D)
    p = O.make_params(shapes, seed, *[float(v) for v in g["gates"]])
    x, xa, mask, _ = O.make_inputs(B, Lq, max(Lk, 1), D, seed, list(g["lengths"]))
    mod = G.GatedCrossModalFusion(D, H, nl, dropout=0.0) if kind == "fusion" else G.GatedCrossAttentionBlock(D, H, dropout=0.0)
    mod.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    return g, kind, (D, H, nl), p, mod.eval(), torch.from_numpy(x), torch.from_numpy(xa), torch.from_numpy(mask)


def _rel(a, b):
    den = float(b.abs().max())
    return float((a.double() - b.double()).abs().max()) / (den if den > 0 else 1.0)


@pytest.mark.parametrize("name", ["block_small", "block_masked", "block_noxa", "fusion_small"])
def test_emitted_graph_reproduces_the_reference(name, stub_forward):
    g, kind, (D, H, nl), p, mod, x, xa, mask = _load(name)
    inputs = (x,) if kind == "noxa" else (x, xa, mask)
    graph, params = E.onnx_graph(mod, inputs)
    hist = onnx_interp.op_histogram(graph)
    assert all(k.startswith("onnx::") and k[6:] in onnx_interp.ALLOWED_OPS for k in hist), hist
    assert hist.get("onnx::MatMul", 0) >= (2 if kind == "noxa" else 6)
    names = [i.debugName() for i in graph.inputs()][:len(inputs)]
    assert names == (["audio_features", "video_features", "video_mask"] if kind == "fusion" else ["x", "xa", "mask"])[:len(inputs)]
    (y,) = onnx_interp.run(graph, params, dict(zip(names, inputs)))
    assert _rel(y, torch.from_numpy(g["y"])) < 2e-6                      # golden output of the live reference module

    # dynamic axes: the same graph at another batch size and other sequence lengths, against the torch port
    B2, Lq2, Lk2 = 3, 11, 19
    x2, xa2, mask2, _ = O.make_inputs(B2, Lq2, Lk2, D, 77, [19, 0, 7])
    x2, xa2, mask2 = torch.from_numpy(x2), torch.from_numpy(xa2), torch.from_numpy(mask2)
    pt = TP.to_torch(p)
    if kind == "fusion":
        ref = TP.fusion_forward(pt, x2, xa2, mask2, H, nl)
    elif kind == "noxa":
        ref = TP.block_forward(pt, x2, None, None, H)
    else:
        ref = TP.block_forward(pt, x2, xa2, mask2, H)
    ref = ref.detach()
    (y2,) = onnx_interp.run(graph, params, dict(zip(names, (x2, xa2, mask2)[:len(inputs)])))
    ok_rows = torch.isfinite(ref).all(dim=-1).all(dim=-1)               # torch's SDPA yields NaN for a sample with no keys
    assert torch.isfinite(y2).all()
    assert _rel(y2[ok_rows], ref[ok_rows]) < 1e-5                        # fp32 port vs float64 graph evaluation


def test_export_accepts_the_4d_memory_and_3d_mask_forms(stub_forward):
    g, kind, (D, H, nl), p, mod, x, xa, mask = _load("block_masked")
    graph, params = E.onnx_graph(mod, (x, xa.unsqueeze(1), mask.unsqueeze(1)))          # reference :35-38
    (y,) = onnx_interp.run(graph, params, {"x": x, "xa": xa.unsqueeze(1), "mask": mask.unsqueeze(1)})
    assert _rel(y, torch.from_numpy(g["y"])) < 2e-6


def test_export_onnx_reports_missing_serialiser(stub_forward, tmp_path):
    g, kind, _, p, mod, x, xa, mask = _load("block_small")
    try:
        import onnx  # noqa: F401
    except ImportError:
        with pytest.raises(RuntimeError, match="onnx"):
            E.export_onnx(mod, str(tmp_path / "block.onnx"), (x, xa, mask))
        return
    E.export_onnx(mod, str(tmp_path / "block.onnx"), (x, xa, mask))
    assert (tmp_path / "block.onnx").stat().st_size > 0


def test_export_rejects_other_modules():
    with pytest.raises(TypeError):
        E.onnx_graph(torch.nn.Linear(4, 4), (torch.zeros(1, 4),))
