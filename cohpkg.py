"""Import helper: the package directory is named ``coherent-rtlsdr_b200`` (hyphen), which the
import statement cannot spell.  ``load()`` returns it as module ``coherent_rtlsdr_b200``."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
_NAME = "coherent_rtlsdr_b200"


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    pkg_dir = os.path.join(ROOT, "coherent-rtlsdr_b200")
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


def synth():
    load()
    import importlib
    return importlib.import_module(_NAME + ".synth")
