"""CPU: the C-ABI library builds, loads and exports every symbol include/dmpc_b200.h declares; struct layouts of the
ctypes mirror match the header (checked through the oracle, which is compiled against the same header)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT
from colaborativempc_b200 import _abi


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "dmpc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dmpc_[a-z_]+)\s*\(", src)))


def test_header_functions_are_exported(built):
    lib = _abi.load_library()
    names = _declared_functions()
    assert len(names) >= 11
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_abi.EXPORTS) == names


def test_struct_layout_matches_header(built):
    from oracle import lpv_oracle as lo
    assert lo.lib().oracle_sizeof_params() == C.sizeof(_abi.DmpcParams)
    # dmpc_options: 3 doubles + 6 int32; dmpc_batch_io: 2 int32 + 15 pointers
    assert C.sizeof(_abi.DmpcOptions) == 3 * 8 + 8 * 4  # 7 ints, padded to the doubles' alignment
    assert C.sizeof(_abi.DmpcBatchIO) == 8 + 15 * 8


def test_default_options_and_version(built):
    lib = _abi.load_library()
    o = _abi.DmpcOptions()
    lib.dmpc_default_options(C.byref(o))
    assert o.polish_delta > 0 and o.cert_tol > 0 and o.polish_refine >= 1 and o.ipm_max_iter >= 10
    assert lib.dmpc_version() >= 100


def test_missing_library_is_loud(tmp_path):
    with pytest.raises(ImportError):
        _abi.load_library(str(tmp_path / "nope.so"))


def test_no_oracle_on_product_path():
    """The product package never imports, links or calls oracle/."""
    pkg = os.path.join(ROOT, "colaborativempc_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("no oracle", ""), os.path.join(dp, f)


def test_create_without_gpu_fails_loudly(built, cfg, track):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from colaborativempc_b200.planner import BatchedPlannerLPV, DmpcError
    from conftest import planner_args
    with pytest.raises(DmpcError):
        BatchedPlannerLPV(*planner_args(cfg, track, 20), nn=2, max_batch=4)
