"""Import shim: exposes tools/make_golden.py (the independent restatement) to the tests."""
import importlib.util
import os

_p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "make_golden.py")
_s = importlib.util.spec_from_file_location("make_golden", _p)
_m = importlib.util.module_from_spec(_s)
_s.loader.exec_module(_m)
globals().update({k: getattr(_m, k) for k in dir(_m) if not k.startswith("_")})
