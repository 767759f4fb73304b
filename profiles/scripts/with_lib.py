"""Run a script of this repo against a measurement build of the library (csrc/Makefile `variant` / `trace` targets):
`python profiles/scripts/with_lib.py <name> <script.py> [args...]` loads lib/libfunasr_b200_<name>.so instead of the product
library.  Profiling only; the product always loads lib/libfunasr_b200.so."""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
from funasr_b200 import _lib  # noqa: E402

name, script = sys.argv[1], sys.argv[2]
_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), f"libfunasr_b200_{name}.so")
assert os.path.exists(_lib.LIB_PATH), _lib.LIB_PATH
sys.argv = [script] + sys.argv[3:]
runpy.run_path(script, run_name="__main__")
