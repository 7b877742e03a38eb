"""Importable alias for the hyphenated package directory ``multimodal-transformer-rl_b200/``.

``import mtrl_b200`` exposes the same modules (``mtrl_b200.model``, ``mtrl_b200._lib`` ...): this
package's ``__path__`` simply points at the product directory, so there is exactly one copy of the code.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
PACKAGE_DIR = _os.path.join(_ROOT, "multimodal-transformer-rl_b200")
__path__.insert(0, PACKAGE_DIR)  # type: ignore[name-defined]

with open(_os.path.join(PACKAGE_DIR, "__init__.py")) as _fh:
    exec(compile(_fh.read(), _os.path.join(PACKAGE_DIR, "__init__.py"), "exec"))
