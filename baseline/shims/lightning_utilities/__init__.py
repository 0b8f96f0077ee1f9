"""Stand-in for ``lightning_utilities`` (only ``compare_version`` / ``package_available``), needed to
import the unmodified reference (``thunder/executors/torch_compile.py:6``). Baseline infrastructure only."""

from lightning_utilities.core.imports import compare_version, package_available  # noqa: F401
