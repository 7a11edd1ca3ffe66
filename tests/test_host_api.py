"""CPU-side tests: the host mirror of the R front end, and that the C-ABI library loads and exports every
symbol include/rootsolve_b200.h declares.  No compute call needs a GPU here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "rootsolve_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rs_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    names = declared_functions()
    assert "rs_stsparse" in names and "rs_convergence_test_noalloc" in names and len(names) >= 25
    for nm in names:
        assert hasattr(built, nm), f"{nm} is declared in rootsolve_b200.h but not exported"
    assert built.rs_abi_version() == 1


def test_header_is_plain_c(tmp_path):
    """the drop-in boundary is a C ABI: include/rootsolve_b200.h must compile as C99 on its own (no C++, no CUDA, no torch
    types), and a C program must link against the library"""
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "t.c"
    src.write_text('#include "rootsolve_b200.h"\n'
                   'int main(void) { rs_stsparse_args a; rs_stsparse_result r; (void)a; (void)r;\n'
                   '                 return rs_abi_version() == 1 && rs_last_error() != 0 ? 0 : 1; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(root, "include"),
                    "-fsyntax-only", str(src)], check=True)
    # the R-side binding shown in INTEGRATION.md (integration/call_stsparse_b200.c), against stand-in R declarations
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(root, "include"),
                    "-I", os.path.join(root, "integration", "mock_r"), "-fsyntax-only",
                    os.path.join(root, "integration", "call_stsparse_b200.c")], check=True)


def build_c_example(tmp_path):
    """compiles and links integration/example_steady3d.c against the in-tree libraries; returns the executable"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "rootsolve_b200")
    exe = str(tmp_path / "example_steady3d")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(root, "include"),
                    os.path.join(root, "integration", "example_steady3d.c"), "-L", libdir, "-lrootsolve_b200", "-lrs_models",
                    "-Wl,-rpath," + libdir, "-o", exe], check=True)
    return exe


def build_user_model(tmp_path):
    """compiles integration/example_model.cu into a model library of its own, the way its header says"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    so = str(tmp_path / "libfisher.so")
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false", "-shared", "-Xcompiler", "-fPIC",
                    "-I", os.path.join(root, "include"), os.path.join(root, "integration", "example_model.cu"), "-o", so], check=True)
    return so


def test_user_model_library_builds_and_exports_its_descriptor(built, tmp_path):
    """a model outside the bundled library (the GPU analogue of a user's DLL): builds from the SDK header alone and is
    found through dllname, as getNativeSymbolInfo(func, PACKAGE = dllname) finds a DLL's function (R/stodes.R:212-220)"""
    import shutil
    from rootsolve_b200 import _capi
    if shutil.which("nvcc") is None:
        pytest.skip("no nvcc")
    so = build_user_model(tmp_path)
    assert _capi.model_symbol("fisher1d", so)
    with pytest.raises(ValueError, match="dyn function not loaded"):
        _capi.model_symbol("diff3d", so)


def test_c_caller_links_and_fails_loudly_without_a_gpu(built, tmp_path):
    """a plain C program against the C ABI: links; and where there is no CUDA device it must stop with the library's
    'no CPU path' error, not compute anything"""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    exe = build_c_example(tmp_path)
    if built.rs_device_count() > 0:
        pytest.skip("a CUDA device is present (the gpu suite runs the example)")
    p = subprocess.run([exe, "8"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert p.returncode == 1 and "no CPU path" in p.stderr


def test_model_library_descriptors(built):
    from rootsolve_b200 import _capi

    class Desc(C.Structure):
        _fields_ = [("abi", C.c_int), ("name", C.c_char_p), ("nparms", C.c_int), ("launch", C.c_void_p),
                    ("initfunc", C.c_void_p), ("initforc", C.c_void_p), ("launch_groups", C.c_void_p)]
    for name, npar in [("diff2d", 7), ("diff3d", 8), ("rd3d", 10), ("rt1d", 7), ("anoxmod", 9), ("diffmap", 8), ("anoxforc", 9)]:
        d = C.cast(_capi.model_symbol(name), C.POINTER(Desc)).contents
        assert d.abi == 2 and d.name.decode() == name and d.nparms == npar and d.launch and d.initfunc
        assert bool(d.initforc) == (name == "anoxforc")
        # the grid models are in accessor form (perturbation on the fly), the small general ones in pointer form
        assert bool(d.launch_groups) == (name in ("diff2d", "diff3d", "rd3d", "rt1d"))
    with pytest.raises(ValueError, match="dyn function not loaded"):
        _capi.model_symbol("no_such_model")


def test_no_cpu_fallback(built):
    """without a GPU every compute entry point fails loudly with RS_ENODEVICE"""
    if built.rs_device_count() > 0:
        pytest.skip("a CUDA device is present")
    import rootsolve_b200 as rs
    with pytest.raises(rs.RootSolveError) as e:
        rs.steady_2D(np.full((5, 5), 10.0), func="diff2d", parms=H.P_DIFF2D_LIN, dimens=[5, 5], ipar=[1, 5, 5, 1, 0])
    assert e.value.code == -2 and "no CPU path" in str(e.value)
    ian = np.array([1, 2], dtype=np.int32)
    jan = np.array([1], dtype=np.int32)
    ngp = C.c_int(0)
    assert built.rs_jgroup(1, H.ip(ian), H.ip(jan), 2, None, C.byref(ngp), None, None) == -2


@pytest.mark.parametrize("st,nnz,n", [("1D", [5, 200, 0], 1000), ("1D", [2, 30, 1], 60), ("2D", [1, 50, 50, 0, 0], 2500),
                                      ("2D", [2, 17, 9, 1, 0], 306), ("2D", [1, 100, 50, 1, 1], 5000),
                                      ("3D", [1, 10, 10, 10, 0, 0, 0], 1000), ("3D", [2, 4, 5, 6, 1, 0, 1], 240)])
def test_sparse_args_match_independent_restatement(oracle, st, nnz, n):
    """Type / nnz vector / ngp / lrw of R/stodes.R:101-160: host mirror vs the oracle's own restatement"""
    from rootsolve_b200.stodes import sparse_args
    Type, v, ngp, lrw, ian, jan = sparse_args(n, st, nnz)
    rType, rv, rngp, rlrw, _ = oracle.stodes_args(n, st, nnz)
    assert (Type, ngp, lrw) == (rType, rngp, rlrw) and np.array_equal(v, rv) and ian is None and jan is None


def test_sparse_args_known_values():
    from rootsolve_b200.stodes import sparse_args
    # the verbose print-out of the reference's 10^3 example: "6910" nonzeros, "1001" function calls (pdf p.16)
    Type, v, ngp, lrw, _, _ = sparse_args(1000, "3D", [1, 10, 10, 10, 0, 0, 0])
    assert (Type, int(v[0]), ngp, lrw) == (4, 6910, 1001, 3 * 1000 + 4 * 6910)
    Type, v, ngp, _, _, _ = sparse_args(1000, "1D", [5, 200, 1])
    assert (Type, int(v[0]), ngp) == (2, 1000 * 7 - 10 + 10, 17)
    # sparseusr: (row, col) pairs sorted by column -> ian/jan
    inz = np.array([[1, 1], [2, 1], [1, 2], [2, 2], [3, 3]])
    Type, v, _, _, ian, jan = sparse_args(3, "sparseusr", inz=inz)
    assert Type == 0 and list(ian) == [1, 3, 5, 6] and list(jan) == [1, 2, 1, 2, 3] and int(v[0]) == 5
    with pytest.raises(ValueError, match="should be sorted"):
        sparse_args(3, "sparseusr", inz=inz[::-1])
    with pytest.raises(ValueError, match="sparsetype not known"):
        sparse_args(3, "4D", [1])


def test_sparse_args_mapped_grids(oracle):
    """2Dmap/3Dmap (R/stodes.R:111-141): Type 30/40, nnz estimate for the FULL grid, the map carried behind TotN"""
    from rootsolve_b200.stodes import sparse_args
    pres, inv = H.make_map(H.ball_mask((6, 7)))
    nnz = [1, 6, 7, 0, 1, len(pres)] + pres.tolist()
    Type, v, ngp, lrw, _, _ = sparse_args(len(inv), "2Dmap", nnz)
    assert Type == 30 and int(v[0]) == 42 * 5 - 2 * 13 + 2 * 7 and ngp == len(inv) + 2
    assert list(v[1:7]) == [1, 6, 7, 0, 1, 42] and np.array_equal(v[7:], pres)
    rType, rv, rngp, rlrw, _ = oracle.stodes_args(len(inv), "2Dmap", nnz)
    assert (Type, ngp, lrw) == (rType, rngp, rlrw) and np.array_equal(v, rv)
    pres3, inv3 = H.make_map(H.ball_mask((4, 5, 6)))
    nnz3 = [1, 4, 5, 6, 0, 0, 0, len(pres3)] + pres3.tolist()
    Type, v, ngp, lrw, _, _ = sparse_args(len(inv3), "3Dmap", nnz3)
    assert Type == 40 and int(v[0]) == 120 * 7 - 3 * 15 and int(v[8]) == 120
    rType, rv, rngp, rlrw, _ = oracle.stodes_args(len(inv3), "3Dmap", nnz3)
    assert (Type, ngp, lrw) == (rType, rngp, rlrw) and np.array_equal(v, rv)


def test_forcing_values_follow_r_approx():
    """forcings (R/stodes.R:229-236): a vector is used as is, a list of (time, value) data sets goes through approx()"""
    from rootsolve_b200.stodes import approx, forcing_values
    ds = np.array([[0.0, 10.0], [10.0, 20.0], [20.0, 0.0]])
    assert approx(ds, 5.0) == 15.0 and approx(ds, 10.0) == 20.0 and approx(ds, 12.5) == 15.0
    assert np.isnan(approx(ds, -1.0)) and np.isnan(approx(ds, 21.0))               # rule = 1
    assert approx(ds, -1.0, rule=2) == 10.0 and approx(ds, 99.0, rule=2) == 0.0
    assert approx(ds, 99.0, rule=(1, 2)) == 0.0 and np.isnan(approx(ds, -1.0, rule=(1, 2)))
    assert approx(ds, 5.0, method="constant") == 10.0 and approx(ds, 5.0, method="constant", f=1.0) == 20.0
    assert approx(ds, 5.0, method="constant", f=0.25) == 12.5
    assert approx(ds[::-1], 5.0) == 15.0                                            # unsorted input
    tied = np.array([[0.0, 1.0], [1.0, 2.0], [1.0, 4.0], [2.0, 5.0]])
    assert approx(tied, 1.0) == 3.0 and approx(tied, 1.0, ties="ordered") == 4.0 and approx(tied, 1.0, ties="min") == 2.0
    assert np.array_equal(forcing_values([100.0, 7.0], 3.0), [100.0, 7.0])
    assert np.array_equal(forcing_values([ds, np.column_stack([ds[:, 0], 2.0 * ds[:, 1]])], 5.0), [15.0, 30.0])
    assert np.array_equal(forcing_values([ds], 25.0, {"rule": 2}), [0.0])
    with pytest.raises(ValueError, match="two-columned"):
        approx(np.zeros((3, 3)), 0.0)


def test_argument_checks_mirror_the_r_front_end(built):
    import rootsolve_b200 as rs
    y = np.full(25, 1.0)
    kw = dict(func="diff2d", parms=H.P_DIFF2D_LIN, ipar=[1, 5, 5, 1, 0])
    with pytest.raises(ValueError, match="dimens should be specified"):
        rs.steady_2D(y, dimens=None, **kw)
    with pytest.raises(ValueError, match="should contain 2 values"):
        rs.steady_2D(y, dimens=[5, 5, 1], **kw)
    with pytest.raises(ValueError, match="not an integer fraction"):
        rs.steady_2D(y, dimens=[4, 4], **kw)
    with pytest.raises(ValueError, match="should contain 3 values"):
        rs.steady_3D(y, dimens=[5, 5], **kw)
    with pytest.raises(ValueError, match="either nspec or dimens"):
        rs.steady_1D(y, **kw)
    with pytest.raises(ValueError, match="cyclicBnd should be a vector or number not exceeding 2"):
        rs.steady_2D(y, dimens=[5, 5], cyclicBnd=3, **kw)
    with pytest.raises(ValueError, match="`atol' must either be a scalar"):
        rs.stodes(y, sparsetype="2D", nnz=[1, 5, 5, 0, 0], atol=[1e-8, 1e-8], **kw)
    with pytest.raises(ValueError, match="maxiter must be >=1"):
        rs.stodes(y, sparsetype="2D", nnz=[1, 5, 5, 0, 0], maxiter=0, **kw)
    with pytest.raises(ValueError, match="should be < the number of state variables"):
        rs.stodes(y, sparsetype="2D", nnz=[1, 5, 5, 0, 0], positive=[1, 26], **kw)
    with pytest.raises(TypeError, match="compiled device model"):
        rs.stodes(y, func=lambda t, y, p: y, sparsetype="2D", nnz=[1, 5, 5, 0, 0])
    with pytest.raises(ValueError, match="must be one of `yale', `ilut' or `ilutp'"):
        rs.stodes(y, sparsetype="2D", nnz=[1, 5, 5, 0, 0], spmethod="gmres", **kw)
    with pytest.raises(ValueError, match="unknown control option"):
        rs.stodes(y, sparsetype="2D", nnz=[1, 5, 5, 0, 0], control={"droptoll": 1e-3}, **kw)
    with pytest.raises(NotImplementedError):
        rs.steady_2D(y, dimens=[5, 5], method="runsteady", **kw)


def test_pack_args_layout(built):
    """the POD block handed to rs_stsparse: R's column-major flattening, itol, Pos/positive, rev(dimens)"""
    import importlib
    host = importlib.import_module("rootsolve_b200.stodes")
    y = np.arange(6.0).reshape(2, 3)           # R: matrix(nrow=2, ncol=3)
    P = host.pack_args(y, 0.0, "diff2d", H.P_DIFF2D_LIN, [1e-6] * 6, 1e-8, 1e-8, "2D", False, [1, 3, 2, 0, 0], None,
                       None, None, [1, 2], 100, None, "rs_models", H.P_DIFF2D_LIN, None, [1, 3, 2, 1, 0], 0, "yale",
                       {"precond": "jacobi", "lin_rtol": 1e-12})
    a = P.args
    assert [a.y[i] for i in range(6)] == [0.0, 3.0, 1.0, 4.0, 2.0, 5.0]
    assert a.neq == 6 and a.itol == 3 and a.Pos == 0 and a.npositive == 2 and a.Type == 3 and a.imethod == 1
    assert [a.nnz[i] for i in range(a.nnnz)] == [6 * 5 - 2 * (3 + 2), 1, 3, 2, 0, 0]
    assert a.ngp == 7 and a.control.contents.precond == 1 and a.control.contents.lin_rtol == 1e-12


def test_slab_bounds(built):
    for d1, nr in [(256, 8), (10, 3), (7, 7), (5, 8), (128, 2)]:
        j0, j1 = C.c_int(0), C.c_int(0)
        edges = []
        for r in range(nr):
            built.rs_slab_bounds(d1, nr, r, C.byref(j0), C.byref(j1))
            edges.append((j0.value, j1.value))
        assert edges[0][0] == 0 and edges[-1][1] == d1
        assert all(edges[r][1] == edges[r + 1][0] for r in range(nr - 1))
        sizes = [b - a for a, b in edges]
        assert max(sizes) - min(sizes) <= 1
