"""Importable alias of the `vllm-v0.1.0_b200/` package directory.

`vllm-v0.1.0_b200` cannot be an import name, so this stub points its own
package path at that directory and executes its __init__.
"""
import os as _os

_impl = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "vllm-v0.1.0_b200")
__path__ = [_impl]
with open(_os.path.join(_impl, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_impl, "__init__.py"), "exec"))
del _f
