"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on identical seeded inputs.

Bars (BASELINE.json north_star): stencil apply / restriction / prolongation <= 1e-13 relative.  The diagonal-metric
kernels are compiled with -fmad=false and follow the reference's statement order, so they are held to BIT-EXACT
equality with the (non-contracting) oracle; full-metric kernels and order-dependent sums to 1e-13.
"""
import numpy as np
import pytest

from common import Problem, rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-13

CASES = {
    "stretched3d": dict(n=(32, 24, 16), map_name="stretched", aniso=4.0),
    "stretched3d_boxes": dict(n=(32, 32, 16), map_name="stretched", aniso=4.0, parts=(1, 4, 1)),
    "periodic_y": dict(n=(32, 16, 16), map_name="stretched", aniso=2.0, periodic=(0, 1, 0)),
    "periodic_xy": dict(n=(16, 16, 16), map_name="cartesian", aniso=1.0, periodic=(1, 1, 0)),
    "dirichlet_top": dict(n=(16, 16, 16), map_name="stretched", aniso=4.0, bc_type=[0, 0, 0, 0, 0, 1],
                          bc_value=[0, 0, 0, 0, 0, 0.5]),
    "helmholtz": dict(n=(16, 16, 16), map_name="stretched", aniso=4.0, alpha=1.0, beta=-0.05),
    "stretched2d": dict(n=(32, 16, 1), dim=2, map_name="stretched", aniso=4.0),
    "terrain3d": dict(n=(32, 24, 16), map_name="terrain", aniso=4.0),
    "terrain3d_boxes": dict(n=(32, 32, 16), map_name="terrain", aniso=4.0, parts=(1, 4, 1)),
    "terrain3d_boxes_xyz": dict(n=(32, 32, 16), map_name="terrain", aniso=4.0, parts=(2, 2, 2)),
    "terrain2d": dict(n=(32, 16, 1), dim=2, map_name="terrain", aniso=4.0),
    "terrain2d_boxes": dict(n=(32, 16, 1), dim=2, map_name="terrain", aniso=4.0, parts=(2, 2, 1)),
    "twisted2d_periodic": dict(n=(32, 32, 1), dim=2, map_name="twisted", aniso=1.0, periodic=(1, 1, 0),
                               map_params={"pert": 0.05}),
    "terrain_periodic_y": dict(n=(16, 16, 16), map_name="terrain", aniso=4.0, periodic=(0, 1, 0)),
}
DIAG = [k for k, v in CASES.items() if v["map_name"] in ("stretched", "cartesian")]
FULL = [k for k in CASES if k not in DIAG]


def _setup(name, oracle_lib, somar, **params):
    pb = Problem(**CASES[name])
    orc = pb.oracle(oracle_lib, **params)
    lev = pb.level(somar, **params)
    return pb, orc, lev


def _check(got, want, exact, what):
    if exact:
        assert np.array_equal(got, want), f"{what}: not bit-exact, rel err {rel_err(got, want):.3e}"
    else:
        assert rel_err(got, want) <= TOL, f"{what}: rel err {rel_err(got, want):.3e}"


@pytest.mark.parametrize("name", list(CASES))
def test_hierarchy_and_metrics(name, oracle_lib, somar):
    pb, orc, lev = _setup(name, oracle_lib, somar)
    assert lev.ndepths == orc.ndepths
    for d in range(lev.ndepths):
        a, b = lev.depth_info(d), orc.depth_info(d)
        assert a["n"] == b["n"] and a["dx"] == b["dx"] and a["zero_avg"] == b["zero_avg"], (d, a, b)
        if d + 1 < lev.ndepths:
            assert a["ratio"] == b["ratio"]
        jg, jinv, lap = lev.metric(d)
        ojg, ojinv, olap = orc.metric(d)
        assert np.array_equal(jinv, ojinv)
        _check(lap, olap, True, f"lapDiag depth {d}")
        for dd in range(pb.dim):
            for c in range(pb.dim):
                if pb.is_diagonal and c != dd:
                    continue
                assert np.array_equal(jg[dd][c], ojg[dd][c]), (d, dd, c)


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("homog", [True, False])
def test_apply_and_residual(name, homog, oracle_lib, somar):
    pb, orc, lev = _setup(name, oracle_lib, somar)
    exact = name in DIAG
    for d in range(lev.ndepths):
        shp = orc.shape(d)
        phi, rhs = pb.rand(shp), pb.rand(shp)
        fphi, frhs, fout = lev.field(d, phi), lev.field(d, rhs), lev.field(d)
        lev.apply_op(fout, fphi, homogeneous=homog)
        _check(fout.get(), orc.apply_op(phi, depth=d, homogeneous=homog), exact, f"applyOp depth {d}")
        lev.residual(fout, fphi, frhs, homogeneous=homog)
        _check(fout.get(), orc.residual(phi, rhs, depth=d, homogeneous=homog), exact, f"residual depth {d}")


def test_relax_line_gsrb_on_vertically_split_boxes_is_rejected(somar):
    """LineGSRB solves per box column in the reference (GSRB.cpp:177-330); the whole-column kernels are that smoother only
    for boxes spanning the vertical, anything else must fail loudly"""
    pb = Problem(**CASES["terrain3d_boxes_xyz"])
    lev = pb.level(somar, relaxMode=3)
    a, b = lev.field(0), lev.field(0)
    with pytest.raises(somar.SomarError):
        lev.relax(a, b, 1)


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("mode", [1, 3, 0])
def test_relax(name, mode, oracle_lib, somar):
    if mode == 3 and CASES[name].get("parts", (1, 1, 1))[CASES[name].get("dim", 3) - 1] > 1:
        pytest.skip("LineGSRB on vertically split boxes is rejected (see the test above)")
    pb, orc, lev = _setup(name, oracle_lib, somar, relaxMode=mode)
    exact = name in DIAG
    for d in range(lev.ndepths):
        shp = orc.shape(d)
        phi, rhs = pb.rand(shp), pb.rand(shp)
        fphi, frhs = lev.field(d, phi), lev.field(d, rhs)
        lev.relax(fphi, frhs, iters=2)
        want = orc.relax(phi, rhs, iters=2, depth=d)
        if exact:
            _check(fphi.get(), want, True, f"relax mode {mode} depth {d}")
        else:
            # two sweeps amplify rounding differences of the cross terms a little
            assert rel_err(fphi.get(), want) <= 10 * TOL, rel_err(fphi.get(), want)

LOOSE_CASES = {
    "boxes_y4": dict(n=(32, 32, 16), map_name="stretched", aniso=4.0, parts=(1, 4, 1)),
    "boxes_xy": dict(n=(32, 32, 16), map_name="stretched", aniso=4.0, parts=(2, 2, 1)),
    "boxes_xyz": dict(n=(32, 32, 16), map_name="stretched", aniso=2.0, parts=(2, 2, 2)),
    "periodic_y_boxes": dict(n=(32, 32, 16), map_name="stretched", aniso=2.0, periodic=(0, 1, 0), parts=(1, 2, 1)),
    "periodic_xy_onebox": dict(n=(16, 16, 16), map_name="cartesian", aniso=1.0, periodic=(1, 1, 0)),
    "dirichlet_top_boxes": dict(n=(16, 32, 16), map_name="stretched", aniso=4.0, bc_type=[0, 0, 0, 0, 0, 1],
                                bc_value=[0, 0, 0, 0, 0, 0.5], parts=(1, 2, 1)),
    "helmholtz_boxes": dict(n=(16, 32, 16), map_name="stretched", aniso=4.0, alpha=1.0, beta=-0.05, parts=(1, 2, 1)),
    "stretched2d_boxes": dict(n=(32, 16, 1), dim=2, map_name="stretched", aniso=4.0, parts=(2, 2, 1)),
}


@pytest.mark.parametrize("name", list(LOOSE_CASES))
def test_relax_loose_gsrb(name, oracle_lib, somar):
    """LooseGSRB (relax mode 2, GSRB.cpp:104-141): box interiors, then box rims with stale cross-box ghosts.  The
    result depends on the box layout, which is the point of the test; bit-exact for diagonal metrics."""
    pb = Problem(**LOOSE_CASES[name])
    orc, lev = pb.oracle(oracle_lib, relaxMode=2), pb.level(somar, relaxMode=2)
    assert lev.relax_kernel(0, 2) == "k_gsrb_loose"
    for d in range(lev.ndepths):
        shp = orc.shape(d)
        phi, rhs = pb.rand(shp), pb.rand(shp)
        fphi, frhs = lev.field(d, phi), lev.field(d, rhs)
        lev.relax(fphi, frhs, iters=3)
        want = orc.relax(phi, rhs, iters=3, depth=d)
        _check(fphi.get(), want, True, f"LooseGSRB depth {d}")
    # it is not LevelGSRB in disguise: with more than one box (or a periodic side) the two smoothers differ
    phi, rhs = pb.rand(orc.shape(0)), pb.rand(orc.shape(0))
    lvl = pb.oracle(oracle_lib, relaxMode=1).relax(phi, rhs, iters=1)
    assert not np.array_equal(lvl, orc.relax(phi, rhs, iters=1))


def test_relax_loose_gsrb_full_metric_is_rejected(somar):
    pb = Problem(n=(16, 16, 8), map_name="terrain", aniso=4.0)
    lev = pb.level(somar, relaxMode=2)
    a, b = lev.field(0), lev.field(0)
    with pytest.raises(somar.SomarError):
        lev.relax(a, b, 1)


@pytest.mark.parametrize("name", list(CASES))
def test_restrict_and_prolong(name, oracle_lib, somar):
    pb, orc, lev = _setup(name, oracle_lib, somar)
    exact = name in DIAG
    for d in range(lev.ndepths - 1):
        shp, shc = orc.shape(d), orc.shape(d + 1)
        phi, rhs, crse = pb.rand(shp), pb.rand(shp), pb.rand(shc)
        fphi, frhs, fc = lev.field(d, phi), lev.field(d, rhs), lev.field(d + 1)
        lev.restrict_residual(fc, fphi, frhs)
        _check(fc.get(), orc.restrict_residual(phi, rhs, depth=d), exact, f"restrictResidual depth {d}")
        fc.set(crse)
        lev.prolong_increment(fphi, fc)
        want = orc.prolong_increment(phi, crse, depth=d)
        # the volume-weighted mean is a sum over all cells: order-dependent, tolerance not bit-exactness
        assert rel_err(fphi.get(), want) <= TOL, rel_err(fphi.get(), want)


@pytest.mark.parametrize("name", ["stretched3d", "periodic_y", "terrain3d"])
def test_blas1_and_norms(name, oracle_lib, somar):
    pb, orc, lev = _setup(name, oracle_lib, somar)
    x, y = pb.rand(), pb.rand()
    fx, fy, fz = lev.field(0, x), lev.field(0, y), lev.field(0)
    assert lev.norm(fx, 0) == np.abs(x).max()
    assert abs(lev.norm(fx, 1) - np.abs(x).sum()) <= 1e-12 * np.abs(x).sum()
    assert abs(lev.norm(fx, 2) - np.sqrt((x * x).sum())) <= 1e-12 * np.sqrt((x * x).sum())
    assert abs(lev.dot(fx, fy) - (x * y).sum()) <= 1e-12 * np.abs(x * y).sum()
    lev.incr(fy, fx, -0.75)
    assert np.array_equal(fy.get(), y + (-0.75) * x)
    lev.axby(fz, fx, fy, 2.0, 3.0)
    assert np.array_equal(fz.get(), 2.0 * x + 3.0 * (y + (-0.75) * x))
    lev.scale(fz, 0.5)
    lev.assign(fy, fz)
    assert np.array_equal(fy.get(), 0.5 * (2.0 * x + 3.0 * (y + (-0.75) * x)))
    lev.set_to_zero(fy)
    assert not fy.get().any()


@pytest.mark.parametrize("name", ["stretched3d", "stretched3d_boxes", "terrain3d"])
def test_fab_round_trip(name, oracle_lib, somar):
    """CHF_FRA path: per-box host FABs with ghost cells in, per-box FABs out."""
    pb, orc, lev = _setup(name, oracle_lib, somar)
    x = pb.rand()
    f = lev.field(0)
    for lo, hi in pb.boxes:
        glo, ghi = [l - 1 for l in lo], [h + 1 for h in hi]
        fab = np.full((ghi[2] - glo[2] + 1, ghi[1] - glo[1] + 1, ghi[0] - glo[0] + 1), np.nan)
        fab[1:-1, 1:-1, 1:-1] = x[lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1]
        f.set_fab(fab, glo, ghi, lo, hi)
    assert np.array_equal(f.get(), x)
    for lo, hi in pb.boxes:
        out = np.zeros((hi[2] - lo[2] + 1, hi[1] - lo[1] + 1, hi[0] - lo[0] + 1))
        f.get_fab(out, lo, hi, lo, hi)
        assert np.array_equal(out, x[lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1])


# --- the three implementations of a LevelGSRB iteration must be bit-identical -----------------------------------
# SOMAR_B200_FUSED = 0: two k_gsrb_ortho colour passes, 1: k_gsrb_fused (register pipeline), 2: k_gsrb_async
# (producer/consumer pipeline over mbarrier rings).  Sizes straddle the 62 x 14 / 60 x 14 core tiles and the
# 64-plane z chunk; BCs exercise the rim path.
VARIANT_CASES = {
    "tile_edges": dict(n=(70, 45, 37), map_name="stretched", aniso=4.0),
    "exact_tile": dict(n=(62, 14, 8), map_name="stretched", aniso=2.0),
    "thin": dict(n=(128, 30, 4), map_name="stretched", aniso=1.0),
    "two_chunks": dict(n=(96, 40, 72), map_name="stretched", aniso=8.0),
    "dirichlet": dict(n=(66, 30, 20), map_name="stretched", aniso=4.0, bc_type=[1, 0, 0, 1, 0, 1],
                      bc_value=[0, 0, 0, 0, 0, 0.5]),
    "helmholtz": dict(n=(64, 32, 16), map_name="stretched", aniso=4.0, alpha=1.0, beta=-0.05),
    # periodic y (the LockExchange 3-D decks): the image rows are ghost rows like a neighbour rank's
    "periodic_y": dict(n=(70, 44, 20), map_name="stretched", aniso=4.0, periodic=(0, 1, 0)),
    "periodic_y_boxes": dict(n=(64, 96, 32), map_name="cartesian", aniso=1.0, periodic=(0, 1, 0), parts=(2, 3, 1)),
}


@pytest.mark.parametrize("name", list(VARIANT_CASES))
def test_gsrb_kernel_variants_bit_identical(name, oracle_lib, somar, monkeypatch):
    pb = Problem(**VARIANT_CASES[name])
    phi, rhs = pb.rand(), pb.rand()
    out = {}
    # "2": the pipelined sweep fed by TMA tensor-map loads (gsrb_tma.cu, the default), "2c": by cp.async (gsrb_async.cu)
    for variant, kernel in (("0", "k_gsrb_ortho"), ("1", "k_gsrb_fused"), ("2", "k_gsrb_tma"), ("2c", "k_gsrb_async")):
        monkeypatch.setenv("SOMAR_B200_FUSED", variant[0])
        monkeypatch.setenv("SOMAR_B200_TMA", "0" if variant == "2c" else "1")
        lev = pb.level(somar, relaxMode=1, maxDepth=0)
        if variant == "1" and any(pb.periodic):
            kernel = "k_gsrb_ortho"   # the register-pipelined predecessor never learnt periodic sides
        assert lev.relax_kernel(0, 1) == kernel
        fphi, frhs = lev.field(0, phi), lev.field(0, rhs)
        lev.relax(fphi, frhs, iters=3)
        out[variant] = fphi.get()
        lev.destroy()
    want = pb.oracle(oracle_lib, relaxMode=1, maxDepth=0).relax(phi, rhs, iters=3, depth=0)
    for variant in out:
        assert np.array_equal(out[variant], want), f"variant {variant}: rel err {rel_err(out[variant], want):.3e}"


LINE_CASES = {
    "odd_nx_dirichlet": dict(n=(67, 30, 20), map_name="stretched", aniso=4.0, bc_type=[1, 0, 0, 1, 0, 1],
                             bc_value=[0, 0, 0, 0, 0, 0.5]),
    "boxes": dict(n=(64, 48, 32), map_name="stretched", aniso=16.0, parts=(1, 4, 1)),
    "helmholtz": dict(n=(34, 18, 16), map_name="stretched", aniso=4.0, alpha=1.0, beta=-0.05),
    "thin": dict(n=(128, 30, 4), map_name="stretched", aniso=1.0),
}


@pytest.mark.parametrize("name", list(LINE_CASES))
def test_line_gsrb_colour_split_layout_bit_identical(name, oracle_lib, somar, monkeypatch):
    """LineGSRB in the colour-split layout (line_compact.cu) against the brick kernel (line.cu) and the oracle: the same
    statements on the same operands, so the same bits -- odd row lengths, Dirichlet x / y sides, alpha != 0 included."""
    pb = Problem(**LINE_CASES[name])
    phi, rhs = pb.rand(), pb.rand()
    out = {}
    for variant, kernel in (("2", "k_line_pass_c"), ("0", "k_line_gsrb")):   # "2": the compact path at every size
        monkeypatch.setenv("SOMAR_B200_LINE_COMPACT", variant)
        lev = pb.level(somar, relaxMode=3, maxDepth=0)
        assert lev.relax_kernel(0, 3) == kernel
        fphi, frhs = lev.field(0, phi), lev.field(0, rhs)
        lev.relax(fphi, frhs, iters=3)
        lev.relax(fphi, frhs, iters=1)            # a second call: phi is packed again, the coefficients are reused
        if name == "helmholtz":
            lev.set_alpha_beta(0.5, -0.1)          # the column matrix changes: coefficients rebuilt
            lev.relax(fphi, frhs, iters=1)
        out[variant] = fphi.get()
        lev.destroy()
    assert np.array_equal(out["2"], out["0"])
    if name != "helmholtz":
        want = pb.oracle(oracle_lib, relaxMode=3, maxDepth=0).relax(phi, rhs, iters=4, depth=0)
        assert np.array_equal(out["2"], want), rel_err(out["2"], want)
