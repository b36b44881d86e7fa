"""SURVEY 8 f4 on the GPU: the export traces the module by running the product's real CUDA forward; the emitted ONNX graph
(evaluated by tests/onnx_interp.py in float64 -- onnxruntime is not in this image) must agree with that forward and with
the golden outputs of the live reference module, at the traced shapes and -- dynamic axes -- at other ones."""
import pytest
import torch

from oracle import gxa_oracle as O
import mocov2_whisper_flamingo_b200 as G
from mocov2_whisper_flamingo_b200 import export as E
import onnx_interp
from test_export import _load, _rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["block_masked", "block_noxa", "fusion_small"])
def test_exported_graph_matches_cuda_forward_and_reference(name):
    g, kind, (D, H, nl), p, mod, x, xa, mask = _load(name)
    mod = mod.cuda()
    G.set_default_precision("fp32")
    try:
        inputs = (x.cuda(),) if kind == "noxa" else (x.cuda(), xa.cuda(), mask.cuda())
        n0 = G.launch_count()
        graph, params = E.onnx_graph(mod, inputs)
        assert G.launch_count() > n0                                      # tracing ran the library's kernels
        names = [i.debugName() for i in graph.inputs()][:len(inputs)]
        (y,) = onnx_interp.run(graph, params, dict(zip(names, inputs)))
        assert _rel(y, torch.from_numpy(g["y"])) < 2e-6                   # reference golden
        with torch.no_grad():
            y_cuda = mod(*inputs) if kind != "fusion" else mod(inputs[0], inputs[1], video_mask=inputs[2])
        assert _rel(y_cuda.cpu(), y) < 1e-4                               # fp32 mode tolerance (north_star)

        B2, Lq2, Lk2 = 3, 11, 19                                          # other batch / lengths through the same graph
        x2, xa2, mask2, _ = O.make_inputs(B2, Lq2, Lk2, D, 77, [19, 0, 7])
        feeds = tuple(torch.from_numpy(t).cuda() for t in (x2, xa2, mask2))[:len(inputs)]
        (y2,) = onnx_interp.run(graph, params, dict(zip(names, feeds)))
        with torch.no_grad():
            y2_cuda = mod(*feeds) if kind != "fusion" else mod(feeds[0], feeds[1], video_mask=feeds[2])
        assert _rel(y2_cuda.cpu(), y2) < 1e-4
    finally:
        G.set_default_precision(None)
