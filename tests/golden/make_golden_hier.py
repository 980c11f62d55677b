"""Writes tests/golden/ref_*.hier / ref_deg*.dhier with the reference's UNMODIFIED HierarchyWriter (oracle/_ref/libref_hier.so =
GH/hierarchy_writer.cpp compiled against oracle/eigen_stub) from the seeded cases of tests/hier_cases.py. Run in the build container
after `make -C oracle ref`:   python tests/golden/make_golden_hier.py"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.dirname(HERE)]
import hier_cases  # noqa: E402
from oracle import refshim_hier as rh  # noqa: E402

h = hier_cases.static_case()
rh.write(os.path.join(HERE, "ref_fp32.hier"), h, False)
rh.write(os.path.join(HERE, "ref_half.hier"), h, True)
for deg in range(4):
    rh.write_dynamic(os.path.join(HERE, "ref_deg%d.dhier" % deg), hier_cases.dynamic_case(deg), deg)
print("written")
