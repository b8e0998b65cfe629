"""C-ABI boundary checks that need no GPU: the library loads, exports every symbol include/lbchem.h declares,
its option defaults are OpenFOAM's, and compute entry points fail loudly (no CPU fallback) without a device."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from loadbalancedchemistrymodel_b200 import capi

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib():
    if not capi.LIB_PATH.exists():
        import __graft_entry__ as g
        g.build()
    return capi.load_library()


def _declared_symbols():
    text = (ROOT / "include/lbchem.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lbchem_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    syms = _declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/lbchem.h but not exported by liblbchem.so"
    assert set(capi.EXPORTED_SYMBOLS) <= set(syms)
    assert set(syms) <= set(capi.EXPORTED_SYMBOLS), "capi.EXPORTED_SYMBOLS is missing header symbols"


def test_version_and_defaults(lib):
    assert lib.lbchem_version().decode().startswith("lbchem")
    o = capi.Options()
    lib.lbchem_default_options(C.byref(o))
    # OpenFOAM ODESolver defaults: absTol SMALL, relTol 1e-4, maxSteps 10000; Treact 0; deltaTChemMax GREAT
    assert o.max_steps == 10000 and o.rel_tol == 1e-4 and o.abs_tol == 1e-15
    assert o.Treact == 0.0 and o.deltaTChemMax == 1e15 and o.cTauChem == 1.0
    assert o.kc_floor == 1e-6 and o.dfdx_is_dcdt == 1
    assert abs(o.gas_constant - capi.RR_DEFAULT) < 1e-9
    p = capi.default_options()
    for name, _ in capi.Options._fields_:
        assert getattr(o, name) == getattr(p, name), name


def test_struct_layouts_match_header(lib):
    """sizeof of the ctypes mirrors equals what the C compiler lays out for include/lbchem.h."""
    import subprocess
    import tempfile
    src = ('#include <stdio.h>\n#include "lbchem.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(lbchem_mech_desc), '
           'sizeof(lbchem_options), sizeof(lbchem_stats), sizeof(lbchem_lb_options), sizeof(lbchem_lb_info), sizeof(lbchem_p2p_op), '
           'sizeof(lbchem_transport));return 0;}\n')
    with tempfile.TemporaryDirectory() as d:
        (Path(d) / "s.c").write_text(src)
        subprocess.check_call(["gcc", "-I", str(ROOT / "include"), str(Path(d) / "s.c"), "-o", str(Path(d) / "s")])
        out = subprocess.check_output([str(Path(d) / "s")]).decode().split()
    assert [int(x) for x in out] == [C.sizeof(capi.MechDesc), C.sizeof(capi.Options), C.sizeof(capi.Stats), C.sizeof(capi.LbOptions),
                                     C.sizeof(capi.LbInfo), C.sizeof(capi.P2POp), C.sizeof(capi.Transport)]


def test_no_cpu_fallback(lib, gri):
    """Without a CUDA device lbchem_create must fail (LBCHEM_E_CUDA), never fall back to host arithmetic."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; this test checks the no-device behaviour")
    with pytest.raises(RuntimeError, match="lbchem_create failed"):
        capi.Chemistry(gri)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.load_library(tmp_path / "liblbchem.so")


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or dlopen it."""
    pkg = ROOT / "loadbalancedchemistrymodel_b200"
    files = []
    for pat in ("*.py", "*.cu", "*.cpp", "*.cuh", "*.hpp", "Makefile"):
        files += list(pkg.rglob(pat))
    files += list((ROOT / "adapter").rglob("*.H")) + list((ROOT / "include").rglob("*.h"))
    for f in files:
        t = f.read_text()
        assert "lbchem_oracle" not in t and "from oracle" not in t and "import oracle" not in t, f


def test_plan_and_cells_to_send_run_on_host(lib):
    send, recv = capi.balance_plan(np.array([4.0, 1.0, 1.0, 2.0]), np.array([10, 10, 10, 10]))
    assert send[0] and not send[1] and not send[2]
    assert capi.cells_to_send(np.ones(10), 3.0, 0) == 3


def test_lb_defaults_and_host_build(lib):
    """LoadBalancedCoeffs defaults (reference LoadBalancedChemistryModel.C:41-45) and the CPU test build of the step."""
    o = capi.LbOptions()
    lib.lbchem_lb_default_options(C.byref(o))
    assert o.update_iter == 0 and o.min_fraction == 0.02 and o.tdac == 0 and o.balance == 1
    host = capi.load_lbhost()
    for sym in ("lbchem_lbhost_create", "lbchem_lb_solve", "lbchem_lb_get_info", "lbchem_lb_get_plan", "lbchem_lb_destroy"):
        assert hasattr(host, sym)
    # NCCL is resolved at run time: the library itself must not depend on libnccl at load time
    import subprocess
    needed = subprocess.check_output(["readelf", "-d", str(capi.LIB_PATH)]).decode()
    assert "libnccl" not in needed
