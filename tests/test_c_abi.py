"""CPU: the C-ABI shared library loads and exports every symbol include/awq_b200.h declares;
without a GPU the compute entry points fail loudly instead of falling back."""
import ctypes
import os
import subprocess

import pytest
import torch

from sglang_awq_b200 import _C


def test_library_built_and_loads():
    assert os.path.exists(_C.LIB_PATH), "run python -m sglang_awq_b200.build"
    assert _C.lib().awq_b200_abi_version() == 1


def test_every_declared_symbol_is_exported():
    declared = _C.declared_symbols()
    assert len(declared) >= 12
    L = ctypes.CDLL(_C.LIB_PATH)
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, f"declared in awq_b200.h but not exported: {missing}"
    assert set(declared) == set(_C._SIGNATURES), "ctypes table and header disagree"


def test_sm100a_sass_is_in_the_library():
    out = subprocess.run(["cuobjdump", "-lelf", _C.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_tile_math():
    L = _C.lib()
    assert L.awq_b200_num_tiles(4096, 11008) == 32 * 86
    assert L.awq_b200_num_tiles(256, 200) == 2 * 2
    assert L.awq_b200_num_tiles(100, 128) == -1
    assert L.awq_b200_workspace_bytes(16, 4096) > 8192 * 4


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    L = _C.lib()
    assert L.awq_b200_check_device() != 0
    assert b"no CPU fallback" in L.awq_b200_last_error()
    q = torch.zeros(128, 16, dtype=torch.int32)
    s = torch.ones(1, 128, dtype=torch.float16)
    out = torch.empty(128, 128, dtype=torch.float16)
    rc = L.awq_b200_dequantize(q.data_ptr(), s.data_ptr(), q.data_ptr(), out.data_ptr(), 128, 128, 1, 0, None)
    assert rc != 0
    rc = L.awq_b200_dequantize(q.data_ptr(), s.data_ptr(), q.data_ptr(), out.data_ptr(), 128, 100, 1, 0, None)
    assert rc != 0 and b"multiple of 8" in L.awq_b200_last_error()


def test_torch_op_schema_matches_reference():
    from sglang_awq_b200 import sgl_kernel  # noqa: F401  (registers the ops)

    schema = str(torch.ops.sgl_kernel.awq_dequantize.default._schema)
    assert schema == "sgl_kernel::awq_dequantize(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor"
    q = torch.zeros(128, 16, dtype=torch.int32, device="meta")
    s = torch.zeros(1, 128, dtype=torch.bfloat16, device="meta")
    out = sgl_kernel.awq_dequantize(q, s, q[:1])
    assert out.shape == (128, 128) and out.dtype == torch.bfloat16
    if not torch.cuda.is_available():
        with pytest.raises(NotImplementedError):
            sgl_kernel.awq_dequantize(torch.zeros(128, 16, dtype=torch.int32), torch.zeros(1, 128, dtype=torch.float16),
                                      torch.zeros(1, 16, dtype=torch.int32))


def test_route_tables_are_consistent():
    """csrc/gemv_route_table.h (generated): the fp16 and the bf16 table cover the same shapes, every rank list names each
    kernel once, and the bf16 one never names the tcgen05 decode kernel (id 5: fp16 only)."""
    import re

    text = open(os.path.join(os.path.dirname(_C.HEADER_PATH), "..", "sglang_awq_b200", "csrc", "gemv_route_table.h")).read()
    tables = {}
    for name in ("kGemvRouteTable", "kGemvRouteTableBf16"):
        body = text[text.index("GemvRouteEntry %s[]" % name):]
        body = body[:body.index("};")]
        rows = re.findall(r"\{(\d+), (\d+), \{(.*)\}\},", body)
        assert len(rows) >= 40
        tables[name] = {}
        for K, N, ranks in rows:
            lists = [[int(v) for v in r.split(",")] for r in re.findall(r"\{([\d,]+)\}", ranks)]
            assert len(lists) == 6 and all(len(r) == 5 for r in lists)
            for r in lists:
                named = [v for v in r if v]
                assert named and len(set(named)) == len(named) and all(1 <= v <= 5 for v in named)
                assert 2 in named, "the legacy kernel takes every problem: it must close every list"
            tables[name][(int(K), int(N))] = lists
    assert set(tables["kGemvRouteTable"]) == set(tables["kGemvRouteTableBf16"])
    assert all(5 not in r for lists in tables["kGemvRouteTableBf16"].values() for r in lists)
