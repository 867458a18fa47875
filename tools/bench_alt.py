"""Run bench.py against an experiment build of the library (make -C rain_b200/csrc variant VARIANT=x EXTRA=-D...):
    python tools/bench_alt.py rain_b200/alt/librainrast_x.so [bench.py arguments]
Tooling for A/B measurements of compile-time choices; the product never loads anything but rain_b200/librainrast.so."""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rain_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.abspath(sys.argv[1])
sys.argv = [os.path.join(ROOT, "bench.py")] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
