"""b200ex — a B200-native fusion executor for Lightning Thunder.

``b200ex.ops``       tensor-level launchers over the C-ABI kernel library (libb200k.so)
``b200ex.executor``  the ``thunder.extend.FusionExecutor`` plugin (imports thunder; import it explicitly)

Importing this package never falls back to another backend: if libb200k.so is absent, the first
launcher call raises ``B200LibraryError``.
"""

from ._lib import B200KernelError, B200LibraryError, lib, library_path  # noqa: F401

__version__ = "0.1.0"
