"""Importable alias of the `nano-vllm-learn_b200/` source tree (a hyphen cannot be imported).

`import nano_vllm_learn_b200` resolves every submodule from `../nano-vllm-learn_b200/`.
"""
import os as _os

_src = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "nano-vllm-learn_b200")
__path__ = [_src]
with open(_os.path.join(_src, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_src, "__init__.py"), "exec"))
del _f
