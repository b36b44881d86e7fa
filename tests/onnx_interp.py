"""A small evaluator for torch's in-memory ONNX graphs (``mocov2_whisper_flamingo_b200.export.onnx_graph``): test
infrastructure that stands in for onnxruntime, which this image does not have.  It implements exactly the opset-14
operators the export emits, with ONNX semantics (Reshape's 0 = "copy this dimension", Slice / Unsqueeze taking their
axes as inputs, numpy broadcasting), on CPU tensors in a chosen floating-point type.
"""
import torch

ALLOWED_OPS = {"Add", "Sub", "Mul", "Div", "Sqrt", "Tanh", "Erf", "ReduceMean", "MatMul", "Transpose", "Reshape", "Slice",
               "Softmax", "Where", "IsNaN", "Unsqueeze", "Squeeze", "Constant", "Identity", "Cast"}


def _attrs(node):
    out = {}
    for a in node.attributeNames():
        kind = node.kindOf(a)                     # "i", "is", "f", "fs", "s", "t", ...
        out[a] = getattr(node, kind)(a)
    return out


def _reshape(x, shape):
    shape = [int(s) for s in shape.tolist()]
    out = [x.shape[i] if s == 0 else s for i, s in enumerate(shape)]
    return x.reshape(out)


def _slice(x, starts, ends, axes=None, steps=None):
    starts, ends = starts.tolist(), ends.tolist()
    axes = list(range(len(starts))) if axes is None else axes.tolist()
    steps = [1] * len(starts) if steps is None else steps.tolist()
    idx = [slice(None)] * x.dim()
    for s, e, a, st in zip(starts, ends, axes, steps):
        idx[a] = slice(int(s), int(e), int(st))
    return x[tuple(idx)]


def _unsqueeze(x, axes):
    for a in sorted(int(a) for a in axes.tolist()):
        x = x.unsqueeze(a)
    return x


def _squeeze(x, axes=None):
    if axes is None:
        return x.squeeze()
    for a in sorted((int(a) for a in axes.tolist()), reverse=True):
        x = x.squeeze(a)
    return x


def _constant(attrs, dtype):
    v = attrs["value"]
    return v.to(dtype) if v.is_floating_point() else v


def run(graph, params, feeds, dtype=torch.float64):
    """Evaluate `graph` on `feeds` (name -> tensor); initialisers come from `params`.  Floating-point values are carried
    in `dtype`.  Returns the list of graph outputs.  Raises on any operator outside ALLOWED_OPS."""
    def conv(t):
        t = t.detach().cpu()
        return t.to(dtype) if t.is_floating_point() else t

    env = {}
    for inp in graph.inputs():
        name = inp.debugName()
        if name in feeds:
            env[name] = conv(feeds[name])
        elif name in params:
            env[name] = conv(params[name])
        else:
            raise KeyError(f"graph input {name} has no value")
    for node in graph.nodes():
        kind = node.kind()
        assert kind.startswith("onnx::"), f"non-ONNX node in the exported graph: {kind}"
        op = kind[6:]
        assert op in ALLOWED_OPS, f"operator outside the documented set: {op}"
        a = _attrs(node)
        x = [env[i.debugName()] if i.debugName() in env else None for i in node.inputs()]
        if op == "Constant":
            y = _constant(a, dtype)
        elif op == "Identity":
            y = x[0]
        elif op == "Cast":
            y = x[0].to({1: dtype, 9: torch.bool, 7: torch.int64, 6: torch.int32, 11: dtype, 10: dtype, 16: dtype}[a["to"]])
        elif op == "Add":
            y = x[0] + x[1]
        elif op == "Sub":
            y = x[0] - x[1]
        elif op == "Mul":
            y = x[0] * x[1]
        elif op == "Div":
            y = x[0] / x[1]
        elif op == "Sqrt":
            y = torch.sqrt(x[0])
        elif op == "Tanh":
            y = torch.tanh(x[0])
        elif op == "Erf":
            y = torch.erf(x[0])
        elif op == "ReduceMean":
            y = x[0].mean(dim=[int(d) for d in a["axes"]], keepdim=bool(a.get("keepdims", 1)))
        elif op == "MatMul":
            y = torch.matmul(x[0], x[1])
        elif op == "Transpose":
            y = x[0].permute(*[int(p) for p in a["perm"]])
        elif op == "Reshape":
            y = _reshape(x[0], x[1])
        elif op == "Slice":
            y = _slice(*x)
        elif op == "Softmax":
            y = torch.softmax(x[0], dim=int(a.get("axis", -1)))
        elif op == "Where":
            y = torch.where(x[0], x[1], x[2])
        elif op == "IsNaN":
            y = torch.isnan(x[0])
        elif op == "Unsqueeze":
            y = _unsqueeze(x[0], x[1])
        elif op == "Squeeze":
            y = _squeeze(*x)
        outs = list(node.outputs())
        assert len(outs) == 1
        env[outs[0].debugName()] = y
    return [env[o.debugName()] for o in graph.outputs()]


def op_histogram(graph):
    h = {}
    for node in graph.nodes():
        h[node.kind()] = h.get(node.kind(), 0) + 1
    return h
