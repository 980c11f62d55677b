"""Drop-in for the reference's `gaussian_hierarchy` package (GH/setup.py): `from gaussian_hierarchy._C import ...`."""
from . import _C  # noqa: F401
from ._C import expand_to_size, expand_to_target, get_interpolation_weights, get_morton_indices, get_spt_cut_cuda, load_dynamic_hierarchy, load_hierarchy, write_dynamic_hierarchy, write_hierarchy  # noqa: F401
