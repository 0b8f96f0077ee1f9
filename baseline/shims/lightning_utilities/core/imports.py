import importlib
import importlib.util
import re

from looseversion import LooseVersion


def package_available(package_name: str) -> bool:
    try:
        return importlib.util.find_spec(package_name) is not None
    except (ModuleNotFoundError, ValueError):
        return False


def compare_version(package: str, op, version: str, use_base_version: bool = False) -> bool:
    try:
        pkg = importlib.import_module(package)
    except ImportError:
        return False
    pkg_version = getattr(pkg, "__version__", None)
    if pkg_version is None:
        return False
    pkg_version = str(pkg_version)
    if use_base_version:
        m = re.match(r"\d+(\.\d+)*", pkg_version)
        pkg_version = m.group(0) if m else pkg_version
    return bool(op(LooseVersion(pkg_version), LooseVersion(version)))
