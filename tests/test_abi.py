"""CPU: the C-ABI library loads, exports every symbol include/lodgs.h declares, and its host-only entry points
(size queries, layouts, argument validation, .hier I/O) behave. No device work is launched here."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import lodgs
from gaussian_hierarchy import _C as gh


def test_exports_every_header_symbol():
    names = lodgs.header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lodgs.lib, n), n
    assert set(names) == set(lodgs._SIGS)


def test_header_cites_reference_and_has_no_torch_types():
    text = open(lodgs.HEADER_PATH).read()
    assert "torch" not in re.sub(r"/\*.*?\*/", "", text, flags=re.S).lower()
    for cite in ("runtime_switching.cu", "rasterizer_impl.cu", "forward.cu", "backward.cu", "hierarchy_loader.cpp", "OurAdam.py"):
        assert cite in text, cite


def test_version_and_arch():
    assert lodgs.lib.lodgs_compiled_sm() == 100
    assert lodgs.lib.lodgs_version() >= 100


def test_buffer_sizes_monotone_and_aligned():
    prev = 0
    for P in (0, 1, 255, 256, 257, 100000, 1000000):
        b = lodgs.lib.lodgs_geom_bytes(P)
        assert b >= prev and b % 128 == 0
        prev = b
    assert lodgs.lib.lodgs_geom_bytes(1000000) < 120 * 1000000  # ~100 B per Gaussian
    assert 1000 * 6 * 4 < lodgs.lib.lodgs_binning_bytes(1000) < 1000 * 6 * 4 + 64 * 1024  # six u32 arrays + sort workspace
    assert lodgs.lib.lodgs_image_bytes(1920, 1080) >= 1920 * 1080 * 8 + 8160 * 8
    o = (C.c_size_t * 9)()
    lodgs.lib.lodgs_geom_layout(1000, o)
    assert all(x % 128 == 0 for x in o) and sorted(o) == list(o)


def test_argument_validation_returns_codes_not_crashes():
    lib = lodgs.lib
    assert lib.lodgs_forward_preprocess(-1, 0, 0, 16, 16, *([None] * 8), 1.0, *([None] * 5), 1.0, 1.0, None, None, 0, None, None) == lodgs.E_ARG
    assert "bad size" in lodgs.last_error()
    assert lib.lodgs_expand_to_size(None, None, 10, 1.0, None, None, None, None, 0, None, 0, 0, None, None) == lodgs.E_ARG
    assert lib.lodgs_sort_pairs(None, None, None, None, None, None, 10, 32, None, 0, None) == lodgs.E_ARG
    assert lib.lodgs_adam_step(None, None, None, None, None, 0, None, 0, 1, 0.1, 0.9, 0.999, 1e-8, 1, 1, None) == lodgs.E_ARG
    with pytest.raises(lodgs.LodgsError):
        lodgs.check(lib.lodgs_hier_probe(b"/nonexistent/file.hier", C.addressof(C.c_int32()), C.addressof(C.c_int32())))


def test_product_has_no_cpu_path():
    from diff_gaussian_rasterization import _C
    e = torch.empty(0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        _C.rasterize_gaussians(torch.zeros(3), e.int(), e.int(), e, e.int(), torch.zeros(4, 3), e, torch.zeros(4, 1), torch.zeros(4, 3),
                               torch.zeros(4, 4), 1.0, e, torch.eye(4), torch.eye(4), 1.0, 1.0, 16, 16, torch.zeros(4, 16, 3), 3,
                               torch.zeros(3), False, False, True)
    with pytest.raises(RuntimeError, match="no CPU path"):
        gh.expand_to_size(torch.zeros(1, 7, dtype=torch.int32), torch.zeros(1, 2, 4), 1.0, torch.zeros(3), torch.zeros(3),
                          torch.zeros(1, dtype=torch.int32), torch.zeros(1, dtype=torch.int32), torch.zeros(1, dtype=torch.int32))


def _rand_hier(P, N, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(P, 3, generator=g), torch.randn(P, 16, 3, generator=g), torch.rand(P, 1, generator=g),
            torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g),
            torch.randint(-1, 30000, (N, 7), generator=g, dtype=torch.int32), torch.randn(N, 2, 4, generator=g))


def test_hier_roundtrip_fp32(tmp_path):
    pos, shs, a, s, r, nodes, boxes = _rand_hier(37, 73)
    f = tmp_path / "a.hier"
    gh.write_hierarchy(str(f), pos, shs, a, s, r, nodes, boxes, False)
    assert os.path.getsize(f) == 4 + 37 * 236 + 4 + 73 * 60
    out = gh.load_hierarchy(str(f))
    for x, y in zip(out, (pos, shs, a, s, r, nodes, boxes)):
        assert x.shape == y.shape and torch.equal(x, y)


def test_hier_roundtrip_half_matches_numpy_float16(tmp_path):
    pos, shs, a, s, r, nodes, boxes = _rand_hier(51, 20, seed=1)
    shs[0, 0, 0], shs[0, 0, 1], shs[0, 0, 2], shs[0, 1, 0] = 65519.0, 1e-8, 6.1e-5, -70000.0   # rounding edge cases
    f = tmp_path / "h.hier"
    gh.write_hierarchy(str(f), pos, shs, a, s, r, nodes, boxes, True)
    assert os.path.getsize(f) == 4 + 51 * 124 + 4 + 20 * 36
    p2, sh2, a2, s2, r2, n2, b2 = gh.load_hierarchy(str(f))
    h = lambda t: torch.from_numpy(t.numpy().astype(np.float16).astype(np.float32))  # noqa: E731
    assert torch.equal(p2, pos) and torch.equal(n2, nodes)
    with np.errstate(over="ignore"):
        for got, want in ((sh2, shs), (a2, a), (s2, s), (r2, r), (b2, boxes)):
            assert torch.equal(got, h(want))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_dynamic_hier_layout_and_roundtrip(tmp_path, deg):
    """.dhier byte layout (GH/hierarchy_writer.cpp:122-168) written by hand with numpy, read by the loader; then writer -> loader."""
    P, M = 29, (deg + 1) ** 2
    g = torch.Generator().manual_seed(deg)
    pos, rot, sca = torch.randn(P, 3, generator=g), torch.randn(P, 4, generator=g), torch.randn(P, 3, generator=g)
    a, shs = torch.rand(P, 1, generator=g), torch.randn(P, M, 3, generator=g)
    nodes = torch.randint(-1, 30000, (P, 6), generator=g, dtype=torch.int32)
    f = tmp_path / "hand.dhier"
    with open(f, "wb") as fh:
        np.array([P, deg], np.int32).tofile(fh)
        for t in (pos, rot, sca, a, shs):
            t.numpy().tofile(fh)
        np.array([P + 5], np.int32).tofile(fh)   # the stored node count is ignored by the reference loader: it reads P nodes
        nodes.numpy().tofile(fh)
    out = gh.load_dynamic_hierarchy(str(f))
    for x, y in zip(out, (pos, shs, a, sca, rot, nodes)):
        assert x.shape == y.shape and x.dtype == y.dtype and torch.equal(x, y)
    f2 = tmp_path / "w.dhier"
    gh.write_dynamic_hierarchy(str(f2), pos, shs, a, sca, rot, nodes, deg)
    assert os.path.getsize(f2) == 8 + P * (12 + 16 + 12 + 4 + 12 * M) + 4 + P * 24
    with open(f, "rb") as fa, open(f2, "rb") as fb:
        ba, bb = bytearray(fa.read()), bytearray(fb.read())
    off = 8 + P * (44 + 12 * M)
    assert ba[:off] == bb[:off] and ba[off + 4:] == bb[off + 4:] and np.frombuffer(bytes(bb[off:off + 4]), np.int32)[0] == P
    with pytest.raises(RuntimeError, match="File not found"):
        gh.load_dynamic_hierarchy("/nonexistent/x.dhier")
    with pytest.raises(RuntimeError, match="shs must hold"):
        gh.write_dynamic_hierarchy(str(f2), pos, shs[:, :1] if deg else torch.zeros(P, 4, 3), a, sca, rot, nodes, deg)
    with open(tmp_path / "short.dhier", "wb") as fh:
        fh.write(bytes(ba[:off - 7]))
    with pytest.raises(RuntimeError, match="truncated"):
        gh.load_dynamic_hierarchy(str(tmp_path / "short.dhier"))


def test_expand_to_target_matches_restated_recursion():
    """Depth-target cut (GH/traversal.cpp:15-39): explicit-stack walk in the library vs the recursion restated in oracle/."""
    from oracle import oracle as orc
    import cases
    nodes, boxes, views = cases.make_cut_case(L=700, seed=3)
    depths = nodes[:, 0]
    for target in (-1, 0, 1, int(np.median(depths)), int(depths.max()), int(depths.max()) + 3):
        mine = gh.expand_to_target(torch.from_numpy(nodes), target)
        want = orc.expand_to_target(nodes, target)
        assert mine.dtype == torch.int32 and np.array_equal(mine.numpy(), want), target
    assert len(orc.expand_to_target(nodes, int(depths.max()) + 3)) == nodes[0, 3] + nodes[0, 4]


def test_import_surface_of_the_reference_scripts():
    """Every name the reference's entry scripts import from the plugin packages resolves (train_single.py:15, train_coarse.py:15,
    train_post.py:17,40, hierarchy_viewer.py:17,40, render_hierarchy.py:19, debug_utils.py:15, scene/gaussian_model.py:24,30,
    utils/reloc_utils.py:1; render_stp exists but raises: its rasterizer is not vendored by the reference)."""
    import importlib
    wanted = {"gaussian_renderer": ["render", "render_post", "render_coarse", "render_on_disk", "render_vanilla", "render_stp",
                                    "network_gui", "occlusion_cull"],
              "diff_gaussian_rasterization": ["GaussianRasterizationSettings", "GaussianRasterizer", "compute_relocation", "_C"],
              "alt_gaussian_rasterization": ["GaussianRasterizationSettings", "GaussianRasterizer", "SparseGaussianAdam", "_C"],
              "gaussian_hierarchy._C": ["load_hierarchy", "load_dynamic_hierarchy", "write_hierarchy", "write_dynamic_hierarchy",
                                        "expand_to_size", "expand_to_size_dynamic", "get_interpolation_weights",
                                        "get_interpolation_weights_dynamic", "expand_to_target", "get_morton_indices", "get_spt_cut_cuda"]}
    for mod, names in wanted.items():
        m = importlib.import_module(mod)
        for n in names:
            assert hasattr(m, n), (mod, n)
    import gaussian_renderer as gr
    import inspect
    assert list(inspect.signature(gr.render_post).parameters)[:16] == [
        "viewpoint_camera", "gaussians", "pipe", "bg_color", "scaling_modifier", "override_color", "render_indices", "parent_indices",
        "interpolation_weights", "num_node_siblings", "interp_python", "on_disk", "use_trained_exp", "iteration", "interp_kernel"][:16]
    assert list(inspect.signature(gr.render_coarse).parameters) == ["viewpoint_camera", "pc", "pipe", "bg_color", "scaling_modifier", "zfar",
                                                                    "override_color", "indices"]
    assert list(inspect.signature(gr.render_vanilla).parameters) == [
        "viewpoint_camera", "means3D", "opacity", "scales", "rotations", "dc", "shs", "pipe", "bg_color", "sh_degree", "scaling_modifier",
        "override_color", "use_trained_exp", "anti_aliasing"]
    assert list(inspect.signature(gr.render_on_disk).parameters) == [
        "viewpoint_camera", "means3D", "opacity", "scales", "rotations", "shs", "pipe", "bg_color", "scaling_modifier", "override_color",
        "sh_degree"]
    with pytest.raises(NotImplementedError):
        gr.render_stp()


def test_load_missing_file_raises_like_reference():
    with pytest.raises(RuntimeError, match="File not found"):
        gh.load_hierarchy("/nonexistent/x.hier")


def test_plugin_surface_matches_reference_names():
    import inspect

    import diff_gaussian_rasterization as d
    import gaussian_renderer as gr
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix", "sh_degree", "campos",
        "prefiltered", "debug", "render_indices", "parent_indices", "interpolation_weights", "num_node_kids", "do_depth")
    assert list(inspect.signature(d.GaussianRasterizer.forward).parameters) == [
        "self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    assert list(inspect.signature(gr.render).parameters) == ["viewpoint_camera", "pc", "pipe", "bg_color", "scaling_modifier",
                                                             "override_color", "indices", "use_trained_exp"]
    assert list(inspect.signature(d._C.rasterize_gaussians).parameters)[:5] == ["background", "indices", "parent_indices", "ts", "kids"]
    assert len(inspect.signature(d._C.rasterize_gaussians).parameters) == 24
    assert len(inspect.signature(d._C.rasterize_gaussians_backward).parameters) == 27
    assert list(inspect.signature(d.compute_relocation).parameters) == ["opacity_old", "scale_old", "N", "binoms", "n_max"]
    # the second plugin (AR/alt_gaussian_rasterization/__init__.py, AR/rasterize_points.cu): 13-field settings, dc + shs, 21 / 27 args
    import alt_gaussian_rasterization as a
    assert a.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix", "sh_degree", "campos",
        "prefiltered", "debug", "antialiasing")
    assert list(inspect.signature(a.GaussianRasterizer.forward).parameters) == [
        "self", "means3D", "means2D", "opacities", "dc", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    assert len(inspect.signature(a._C.rasterize_gaussians).parameters) == 21
    assert len(inspect.signature(a._C.rasterize_gaussians_backward).parameters) == 27
    assert list(inspect.signature(a._C.adamUpdate).parameters) == ["param", "param_grad", "exp_avg", "exp_avg_sq", "visible", "lr", "b1", "b2",
                                                                   "eps", "N", "M"]
    assert list(inspect.signature(a.SparseGaussianAdam.step).parameters) == ["self", "visibility", "N"]
    import gaussian_hierarchy._C as ghc
    assert list(inspect.signature(ghc.expand_to_size_dynamic).parameters) == [
        "nodes", "positions", "scales", "size", "viewpoint", "viewdir", "render_indices", "parent_indices", "nodes_for_render_indices"]
    assert list(inspect.signature(ghc.get_interpolation_weights_dynamic).parameters) == [
        "indices", "size", "nodes", "positions", "scales", "viewpoint", "viewdir", "ts", "num_kids"]
    assert list(inspect.signature(ghc.get_spt_cut_cuda).parameters) == [
        "number_of_SPTs", "gaussian_indices", "SPT_starts", "SPT_max", "SPT_min", "SPT_indices", "SPT_distances"]
    assert list(inspect.signature(ghc.get_morton_indices).parameters) == ["xyz", "min", "max", "code"]
    assert list(inspect.signature(ghc.load_dynamic_hierarchy).parameters) == ["filename"]
    assert list(inspect.signature(ghc.write_dynamic_hierarchy).parameters) == [
        "filename", "pos", "shs", "opacities", "log_scales", "rotations", "nodes", "SH_degree"]
    assert list(inspect.signature(ghc.expand_to_target).parameters) == ["nodes", "target"]
    from lodgs.optim import Adam
    assert list(inspect.signature(Adam.step).parameters)[:2] == ["self", "relevant"]
    r = d.GaussianRasterizer(None)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r.forward(None, None, None)
