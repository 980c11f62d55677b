"""ctypes binding of liblodgs.so (include/lodgs.h). PyTorch is used only for device memory and streams.

The library is mandatory: `lodgs.load()` — called by the first use of `lodgs.lib` and by every plugin module at import — fails loudly
when liblodgs.so is missing or does not export every symbol of the header; there is no CPU or PyTorch fallback anywhere in the product
path. Importing the bare package (for `lodgs.synth`, the seeded input generators) does not map the library, so a process that only
runs the reference's code never has it loaded.
"""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_PKG, "lib", "liblodgs.so")
HEADER_PATH = os.path.join(os.path.dirname(_PKG), "include", "lodgs.h")

OK, E_ARG, E_WORKSPACE, E_CUDA, E_IO = 0, -1, -2, -3, -4
RASTER_SQUARE_RECT, RASTER_NO_ANTIALIAS, RASTER_BACKGROUND_TWICE, RASTER_CLAMPED_ALPHA_GRAD = 1, 2, 4, 8  # LODGS_RASTER_* of include/lodgs.h


class LodgsError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("lodgs error %d: %s" % (code, text))
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("_lodgs_build", os.path.join(_PKG, "build.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build(verbose=False)
        except Exception as e:  # noqa: BLE001
            raise ImportError("liblodgs.so is missing at %s and could not be built: %s" % (LIB_PATH, e))
    return C.CDLL(LIB_PATH)


class _Lib:
    """`lodgs.lib`: binds liblodgs.so on first attribute access."""
    _cdll = None

    def __getattr__(self, name):
        return getattr(load(), name)


lib = _Lib()

_p, _i, _f, _sz, _i64, _d = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64, C.c_double

_SIGS = {
    "lodgs_last_error": (C.c_char_p, []),
    "lodgs_version": (_i, []),
    "lodgs_compiled_sm": (_i, []),
    "lodgs_launch_count": (C.c_uint64, []),
    "lodgs_profile": (_i, [_i]),
    "lodgs_profile_dump": (_i, [C.c_char_p, _sz]),
    "lodgs_cut_workspace_bytes": (_sz, [_i]),
    "lodgs_expand_to_size": (_i, [_p, _p, _i, _f, _p, _p, _p, _p, _i, _p, _sz, _i, _p, _p]),
    "lodgs_cut_prepare": (_i, [_p, _p, _i, _p, _sz, _p, _p]),
    "lodgs_get_interpolation_weights": (_i, [_p, _i, _f, _p, _p, _f, _f, _f, _p, _p, _p]),
    "lodgs_expand_to_size_dynamic": (_i, [_p, _p, _p, _i, _f, _p, _f, _f, _f, _p, _p, _p, _i, _p, _sz, _p, _p]),
    "lodgs_get_interpolation_weights_dynamic": (_i, [_p, _i, _f, _p, _p, _p, _f, _f, _f, _p, _p, _p]),
    "lodgs_hier_probe": (_i, [C.c_char_p, _p, _p]),
    "lodgs_hier_load": (_i, [C.c_char_p, _p, _p, _p, _p, _p, _p, _p]),
    "lodgs_hier_write": (_i, [C.c_char_p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _i]),
    "lodgs_expand_to_target": (_i, [_p, _i, _i, _p, C.c_int64, _p]),
    "lodgs_dhier_probe": (_i, [C.c_char_p, _p, _p, _p]),
    "lodgs_dhier_load": (_i, [C.c_char_p, _p, _p, _p, _p, _p, _p]),
    "lodgs_dhier_write": (_i, [C.c_char_p, _i, _i, _i, _p, _p, _p, _p, _p, _p]),
    "lodgs_hier_build_workspace_bytes": (_sz, [_i]),
    "lodgs_creator_activate": (_i, [_p, _i64, _p, _p, _p, _p, _p, _p]),
    "lodgs_logf_host": (_i, [_p, _i64, _p]),
    "lodgs_hier_build": (_i, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "lodgs_geom_bytes": (_sz, [_i]),
    "lodgs_image_bytes": (_sz, [_i, _i]),
    "lodgs_binning_bytes": (_sz, [_i]),
    "lodgs_forward_preprocess": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                                      _p, _sz, _p, _p]),
    "lodgs_forward_preprocess_ex": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                                         _p, _sz, _p, C.c_uint, _p]),
    "lodgs_forward_preprocess_lod": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                                          _p, _sz, _p, C.c_uint, _p]),
    "lodgs_forward_render": (_i, [_i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _sz, _p, _sz, _p, _p, _p, _p]),
    "lodgs_backward": (_i, [_i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                            _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "lodgs_backward_ex": (_i, [_i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                               _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, C.c_uint, _p]),
    "lodgs_backward_rgb": (_i, [_i, _i, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                                _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, C.c_uint, _p]),
    "lodgs_sh_adam_from_rgb": (_i, [_i64, _i, _i, _p, _i64, _p, _p, _p, _i64, _p, _p, _p, _p, _i, _p, _p, _p, _i, _i, _d, _d, _d, _d, _d,
                                    _i64, _i, _p]),
    "lodgs_ar_bucket_count": (_i, [_i, _i, _i, _i, _p, _p, _p, _sz, _p, _p]),
    "lodgs_mark_visible": (_i, [_i, _p, _p, _p, _p, _p]),
    "lodgs_spt_cut_workspace_bytes": (_sz, [_i]),
    "lodgs_spt_cut_count": (_i, [_i, _p, _i, _p, _p, _p, _p, _p, _p, _p, _sz, _p, _p]),
    "lodgs_spt_cut_write": (_i, [_i, _p, _i, _p, _p, _p, _p, _p, _p, _i, _p]),
    "lodgs_upper_cut_workspace_bytes": (_sz, [_i]),
    "lodgs_upper_tree_cut": (_i, [_p, _i, _i, _p, _p, _p, _i, _p, _p, _f, _p, _i, _p, _p, _p, _p, _p, _sz, _p, _p]),
    "lodgs_page_workspace_bytes": (_sz, [_i64]),
    "lodgs_page_plan": (_i, [_p, _i64, _p, _p, _sz, _p, _p]),
    "lodgs_page_plan_static_max_rows": (_i64, []),
    "lodgs_page_plan_static": (_i, [_p, _i64, _p, _p, _p]),
    "lodgs_page_apply": (_i, [_p, _i64, _i, _p, _p, _p, _p, _i64, _i64, _p]),
    "lodgs_host_scatter_rows": (_i, [_p, _i64, _i64, _p, _i64, _p]),
    "lodgs_host_gather_rows": (_i, [_p, _i64, _i64, _p, _i64, _p]),
    "lodgs_l1_ssim_workspace_bytes": (_sz, [_i, _i, _i]),
    "lodgs_l1_ssim_loss": (_i, [_p, _p, _i, _i, _i, _f, _f, _f, _p, _p, _p, _sz, _p]),
    "lodgs_get_morton_codes": (_i, [_p, _p, _p, _p, _i, _p]),
    "lodgs_compute_relocation": (_i, [_i, _p, _p, _p, _p, _i, _p, _p, _p]),
    "lodgs_geom_layout": (_i, [_i, _p]),
    "lodgs_binning_layout": (_i, [_i, _p]),
    "lodgs_image_layout": (_i, [_i, _i, _p]),
    "lodgs_blend_stats": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "lodgs_sorted_keys": (_i, [_i, _i, _p, _p, _p, _p]),
    "lodgs_sort_workspace_bytes": (_sz, [_i]),
    "lodgs_sort_pairs": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _p, _sz, _p]),
    "lodgs_backward_overwrites_dsh": (_i, [_i, _i, _i, _p, _p]),
    "lodgs_backward_overwrites_geometry": (_i, [_i, _i]),
    "lodgs_adam_step": (_i, [_p, _p, _p, _p, _p, _i64, _p, _i64, _i, _d, _d, _d, _d, _i64, _i, _p]),
    "lodgs_adam_step_multi": (_i, [_i, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p]),
}


def header_symbols():
    """Names of all functions declared in include/lodgs.h."""
    with open(HEADER_PATH) as fh:
        text = fh.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lodgs_[a-z0-9_]+)\s*\(", text)))


def _bind(cdll):
    missing = []
    for name, (res, args) in _SIGS.items():
        try:
            fn = getattr(cdll, name)
        except AttributeError:
            missing.append(name)
            continue
        fn.restype = res
        fn.argtypes = args
    if missing:
        raise ImportError("liblodgs.so does not export: %s" % ", ".join(missing))


def load():
    """Map liblodgs.so and bind every entry point of include/lodgs.h (idempotent). Raises ImportError when that is not possible."""
    if _Lib._cdll is None:
        cdll = _load()
        _bind(cdll)
        _Lib._cdll = cdll
    return _Lib._cdll


def loaded():
    return _Lib._cdll is not None


def last_error():
    return lib.lodgs_last_error().decode()


def check(code):
    if code != 0:
        raise LodgsError(code, last_error())
    return code


def ptr(t):
    """Device/host pointer of a torch tensor; None or an empty tensor -> NULL (the reference's empty-tensor convention)."""
    if t is None:
        return None
    if t.numel() == 0:
        return None
    return t.data_ptr()


def current_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def launch_count():
    return int(lib.lodgs_launch_count())


def profile(enable):
    check(lib.lodgs_profile(1 if enable else 0))


def profile_dump():
    """-> {kernel_name: (total_ms, launches)} since profile(True)."""
    buf = C.create_string_buffer(1 << 16)
    check(lib.lodgs_profile_dump(buf, len(buf)))
    out = {}
    for line in buf.value.decode().splitlines():
        name, ms, n = line.rsplit(" ", 2)
        out[name] = (float(ms), int(n))
    return out
