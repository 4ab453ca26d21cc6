"""Importable alias of the package directory
`project-chronos-distributed-behavioral-edr-ebpf-llm-_b200/` (its name, fixed by the
project layout, is not a Python identifier). `import chronos_b200` loads that package."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "project-chronos-distributed-behavioral-edr-ebpf-llm-_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
