"""Import shim: the package directory is `llm-slo-ebpf-toolkit_b200/` (the name the project layout
prescribes), which Python cannot import because of the hyphens.  This shim makes it importable as
`llm_slo_ebpf_toolkit_b200` by pointing __path__ at that directory and running its __init__."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "llm-slo-ebpf-toolkit_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _f
