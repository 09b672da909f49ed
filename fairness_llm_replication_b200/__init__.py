"""Import alias.  The package source lives in ``../fairness-llm-replication_b200/`` (the
directory name carries the reference's hyphenated name and so is not importable by name);
this shim points ``__path__`` at it and runs its ``__init__``."""
import os as _os

_src = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                     "fairness-llm-replication_b200")
__path__ = [_src]
with open(_os.path.join(_src, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_src, "__init__.py"), "exec"))
del _f
