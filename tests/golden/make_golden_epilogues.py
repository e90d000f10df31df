"""Golden vectors for the fused epilogues either side of the AWQ linears, made by
EXECUTING the reference's own Python (build container only: /root/reference is not on the GPU box):

  RMSNorm.forward_native (with residual)   python/sglang/srt/layers/layernorm.py:178-222
  SiluAndMul.forward_native                python/sglang/srt/layers/activation.py:68-70

The two methods are extracted with `ast` (their modules import the serving runtime) and run unmodified
on CPU tensors against a stand-in `self` that carries the attributes they read.

    python tests/golden/make_golden_epilogues.py      # rewrites tests/golden/epilogue_*.npz
"""
import ast
import os
import types

import numpy as np
import torch
import torch.nn.functional as F

REF = os.environ.get("AWQ_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def extract_method(path, cls, name):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for f in node.body:
                if isinstance(f, ast.FunctionDef) and f.name == name:
                    f.decorator_list = []
                    ns = {"torch": torch, "F": F, "Optional": __import__("typing").Optional,
                          "Union": __import__("typing").Union, "Tuple": __import__("typing").Tuple}
                    exec(compile(ast.Module(body=[f], type_ignores=[]), path, "exec"), ns)
                    return ns[name]
    raise KeyError((cls, name))


def bits(t):
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def main():
    rms = extract_method(os.path.join(REF, "python/sglang/srt/layers/layernorm.py"), "RMSNorm", "forward_native")
    silu = extract_method(os.path.join(REF, "python/sglang/srt/layers/activation.py"), "SiluAndMul", "forward_native")
    g = torch.Generator().manual_seed(20261105)
    for name, rows, hidden, dt, eps in (("f16_h1024", 3, 1024, torch.float16, 1e-6), ("f16_h5120", 2, 5120, torch.float16, 1e-5),
                                        ("bf16_h2048", 4, 2048, torch.bfloat16, 1e-6), ("f16_h200", 5, 200, torch.float16, 1e-6)):
        x = (torch.randn(rows, hidden, generator=g) * 0.7).to(dt)
        res = (torch.randn(rows, hidden, generator=g) * 1.3).to(dt)
        w = (torch.rand(hidden, generator=g) * 1.5 + 0.25).to(dt)
        me = types.SimpleNamespace(weight=w, hidden_size=hidden, variance_epsilon=eps, variance_size_override=None,
                                   cast_x_before_out_mul=False, fp32_residual=False, override_orig_dtype=None)
        out, res_out = rms(me, x.clone(), res.clone())
        assert out.dtype == dt and res_out.dtype == dt
        np.savez_compressed(os.path.join(HERE, f"epilogue_rmsnorm_{name}.npz"), x=bits(x), residual=bits(res), weight=bits(w),
                            out=bits(out), residual_out=bits(res_out), eps=np.float32(eps),
                            dtype=np.int32(0 if dt == torch.float16 else 1))
        print("rmsnorm", name, "ok")
    for name, rows, d, dt in (("f16_d512", 3, 512, torch.float16), ("bf16_d256", 4, 256, torch.bfloat16)):
        x = (torch.randn(rows, 2 * d, generator=g) * 2.0).to(dt)
        out = silu(types.SimpleNamespace(), x)
        assert out.dtype == dt
        np.savez_compressed(os.path.join(HERE, f"epilogue_silu_mul_{name}.npz"), x=bits(x), out=bits(out),
                            dtype=np.int32(0 if dt == torch.float16 else 1))
        print("silu_mul", name, "ok")


if __name__ == "__main__":
    main()
