"""Helpers for the -m gpu tests (torch <-> uint16 bit patterns, synthetic inputs on device)."""
import numpy as np
import torch

TDT = {0: torch.float16, 1: torch.bfloat16}


def to_bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


def from_bits(bits: np.ndarray, dt: int, device="cuda") -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(bits).view(np.int16)).view(TDT[dt]).to(device)


def dev_inputs(oracle, K, N, G, dt, seed=1234, M=0, bias=False, signed_scales=False, device="cuda"):
    inp = oracle.make_inputs(K, N, G, dt, seed=seed, M=M, bias=bias, signed_scales=signed_scales)
    d = {"np": inp,
         "qweight": torch.from_numpy(inp["qweight"]).to(device),
         "qzeros": torch.from_numpy(inp["qzeros"]).to(device),
         "scales": from_bits(inp["scales"], dt, device)}
    if M:
        d["x"] = from_bits(inp["x"], dt, device)
    if bias:
        d["bias"] = from_bits(inp["bias"], dt, device)
    return d


def big_inputs(K, N, G, dtype, M=0, seed=1234, device="cuda"):
    """Device-side generation for full-size shapes (no oracle copy)."""
    g = torch.Generator(device=device).manual_seed(seed)
    qweight = torch.randint(-(2**31), 2**31, (K, N // 8), generator=g, device=device, dtype=torch.int64).to(torch.int32)
    qzeros = torch.randint(-(2**31), 2**31, (K // G, N // 8), generator=g, device=device, dtype=torch.int64).to(torch.int32)
    scales = (torch.rand(K // G, N, generator=g, device=device) * 0.02 + 0.005).to(dtype)
    out = {"qweight": qweight, "qzeros": qzeros, "scales": scales}
    if M:
        out["x"] = torch.randn(M, K, generator=g, device=device).to(dtype)
    return out


def rel_fro(a: torch.Tensor, ref: torch.Tensor) -> float:
    a = a.detach().double()
    ref = ref.double()
    return float(((a - ref).norm() / ref.norm().clamp_min(1e-30)).detach())
