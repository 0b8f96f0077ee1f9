"""Minimal pure-Python stand-in for the ``optree`` package.

The reference (lightning-thunder) imports ``optree`` in ``thunder/core/pytree.py:5`` and it is
not installable in this sandbox (no network).  This module implements exactly the surface the
reference touches (``register_pytree_node``, ``tree_flatten``, ``tree_unflatten``, ``tree_map``,
``tree_iter`` and ``PyTreeSpec.unflatten/num_leaves``) so that the UNMODIFIED reference can be
imported as the parity oracle and as the host of the b200 executor plugin.

It is test/baseline infrastructure, not product code.

Semantics kept from optree: ``None`` is a leaf when ``none_is_leaf=True``; dict keys are visited
in sorted order (insertion order if unsortable); tuples, lists, dicts, OrderedDict, defaultdict,
deque and namedtuples are interior nodes; custom nodes are looked up per namespace.

Public callables are exported as bound ``__call__`` method-wrappers so Thunder's bytecode
interpreter treats them as opaque builtins instead of trying to interpret them
(``thunder/core/interpreter.py:769-779``).
"""

from __future__ import annotations

import collections
from typing import Any, Callable, Iterable

__version__ = "0.0.shim"

_REGISTRY: dict[tuple[str, type], tuple[Callable, Callable]] = {}


def _is_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(type(obj), "_fields") and hasattr(type(obj), "_make")


def _sorted_keys(d):
    try:
        return sorted(d.keys())
    except TypeError:
        return list(d.keys())


class PyTreeSpec:
    __slots__ = ("kind", "node_type", "metadata", "children_specs", "num_leaves", "num_nodes", "none_is_leaf", "namespace", "unflatten_fn")

    def __init__(self, kind, node_type, metadata, children_specs, none_is_leaf, namespace, unflatten_fn=None):
        self.kind = kind  # "leaf" | "none" | "tuple" | "list" | "dict" | "odict" | "ddict" | "deque" | "namedtuple" | "custom"
        self.node_type = node_type
        self.metadata = metadata
        self.children_specs = children_specs
        self.none_is_leaf = none_is_leaf
        self.namespace = namespace
        self.unflatten_fn = unflatten_fn
        if kind == "leaf":
            self.num_leaves = 1
        elif kind == "none":
            self.num_leaves = 0
        else:
            self.num_leaves = sum(c.num_leaves for c in children_specs)
        self.num_nodes = 1 + sum(c.num_nodes for c in children_specs)

    @property
    def type(self):
        """Container type of the root node (``None`` for a leaf), as optree.PyTreeSpec.type."""
        return None if self.kind == "leaf" else (type(None) if self.kind == "none" else self.node_type)

    def children(self):
        return list(self.children_specs)

    def unflatten(self, leaves: Iterable[Any]):
        leaves = list(leaves)
        if len(leaves) != self.num_leaves:
            raise ValueError(f"treespec.unflatten(leaves): expected {self.num_leaves} leaves, got {len(leaves)}.")
        it = iter(leaves)
        return self._build(it)

    def _build(self, it):
        k = self.kind
        if k == "leaf":
            return next(it)
        if k == "none":
            return None
        kids = [c._build(it) for c in self.children_specs]
        if k == "tuple":
            return tuple(kids)
        if k == "list":
            return kids
        if k == "dict":
            return dict(zip(self.metadata, kids))
        if k == "odict":
            return collections.OrderedDict(zip(self.metadata, kids))
        if k == "ddict":
            factory, keys = self.metadata
            return collections.defaultdict(factory, zip(keys, kids))
        if k == "deque":
            return collections.deque(kids, maxlen=self.metadata)
        if k == "namedtuple":
            return self.node_type(*kids)
        if k == "custom":
            return self.unflatten_fn(self.metadata, kids)
        raise AssertionError(k)

    def __eq__(self, other):
        if not isinstance(other, PyTreeSpec):
            return NotImplemented
        return (
            self.kind == other.kind
            and self.node_type is other.node_type
            and self.none_is_leaf == other.none_is_leaf
            and _meta_eq(self.metadata, other.metadata)
            and self.children_specs == other.children_specs
        )

    def __hash__(self):
        return hash((self.kind, self.node_type, self.num_leaves, self.num_nodes))

    def __repr__(self):
        return f"PyTreeSpec({self._fmt()})"

    def _fmt(self):
        k = self.kind
        if k == "leaf":
            return "*"
        if k == "none":
            return "None"
        kids = [c._fmt() for c in self.children_specs]
        if k == "tuple":
            return "(" + ", ".join(kids) + ("," if len(kids) == 1 else "") + ")"
        if k == "list":
            return "[" + ", ".join(kids) + "]"
        if k in ("dict", "odict"):
            return "{" + ", ".join(f"{key!r}: {c}" for key, c in zip(self.metadata, kids)) + "}"
        return f"{getattr(self.node_type, '__name__', self.node_type)}({', '.join(kids)})"


def _meta_eq(a, b):
    try:
        return bool(a == b)
    except Exception:
        return a is b


_LEAF_CACHE: dict[tuple[bool, str], PyTreeSpec] = {}


def _flatten_into(tree, leaves: list, none_is_leaf: bool, namespace: str) -> PyTreeSpec:
    t = type(tree)
    if tree is None:
        if none_is_leaf:
            leaves.append(None)
            return PyTreeSpec("leaf", None, None, (), none_is_leaf, namespace)
        return PyTreeSpec("none", type(None), None, (), none_is_leaf, namespace)

    entry = _REGISTRY.get((namespace, t)) or _REGISTRY.get(("", t))
    if entry is not None:
        flatten_fn, unflatten_fn = entry
        res = flatten_fn(tree)
        children, metadata = res[0], res[1]
        specs = tuple(_flatten_into(c, leaves, none_is_leaf, namespace) for c in children)
        return PyTreeSpec("custom", t, metadata, specs, none_is_leaf, namespace, unflatten_fn)

    if t is tuple:
        specs = tuple(_flatten_into(c, leaves, none_is_leaf, namespace) for c in tree)
        return PyTreeSpec("tuple", tuple, None, specs, none_is_leaf, namespace)
    if t is list:
        specs = tuple(_flatten_into(c, leaves, none_is_leaf, namespace) for c in tree)
        return PyTreeSpec("list", list, None, specs, none_is_leaf, namespace)
    if t is dict:
        keys = _sorted_keys(tree)
        specs = tuple(_flatten_into(tree[k], leaves, none_is_leaf, namespace) for k in keys)
        return PyTreeSpec("dict", dict, keys, specs, none_is_leaf, namespace)
    if t is collections.OrderedDict:
        keys = list(tree.keys())
        specs = tuple(_flatten_into(tree[k], leaves, none_is_leaf, namespace) for k in keys)
        return PyTreeSpec("odict", t, keys, specs, none_is_leaf, namespace)
    if t is collections.defaultdict:
        keys = _sorted_keys(tree)
        specs = tuple(_flatten_into(tree[k], leaves, none_is_leaf, namespace) for k in keys)
        return PyTreeSpec("ddict", t, (tree.default_factory, keys), specs, none_is_leaf, namespace)
    if t is collections.deque:
        specs = tuple(_flatten_into(c, leaves, none_is_leaf, namespace) for c in tree)
        return PyTreeSpec("deque", t, tree.maxlen, specs, none_is_leaf, namespace)
    if _is_namedtuple(tree):
        specs = tuple(_flatten_into(c, leaves, none_is_leaf, namespace) for c in tree)
        return PyTreeSpec("namedtuple", t, None, specs, none_is_leaf, namespace)

    leaves.append(tree)
    return PyTreeSpec("leaf", None, None, (), none_is_leaf, namespace)


def _register_pytree_node(cls, flatten_func, unflatten_func, *, namespace=""):
    if not isinstance(cls, type):
        raise TypeError(f"Expected a class, got {cls!r}.")
    key = (namespace, cls)
    if key in _REGISTRY:
        raise ValueError(f"PyTree type {cls!r} is already registered in namespace {namespace!r}.")
    _REGISTRY[key] = (flatten_func, unflatten_func)
    return cls


def _tree_flatten(tree, is_leaf=None, *, none_is_leaf=False, namespace=""):
    leaves: list = []
    spec = _flatten_into(tree, leaves, none_is_leaf, namespace)
    return leaves, spec


def _tree_unflatten(treespec, leaves):
    return treespec.unflatten(leaves)


def _tree_iter(tree, is_leaf=None, *, none_is_leaf=False, namespace=""):
    leaves, _ = _tree_flatten(tree, none_is_leaf=none_is_leaf, namespace=namespace)
    return iter(leaves)


def _tree_leaves(tree, is_leaf=None, *, none_is_leaf=False, namespace=""):
    return _tree_flatten(tree, none_is_leaf=none_is_leaf, namespace=namespace)[0]


def _tree_structure(tree, is_leaf=None, *, none_is_leaf=False, namespace=""):
    return _tree_flatten(tree, none_is_leaf=none_is_leaf, namespace=namespace)[1]


def _tree_map(func, tree, *rests, is_leaf=None, none_is_leaf=False, namespace=""):
    leaves, spec = _tree_flatten(tree, none_is_leaf=none_is_leaf, namespace=namespace)
    cols = [leaves]
    for r in rests:
        rl, rs = _tree_flatten(r, none_is_leaf=none_is_leaf, namespace=namespace)
        if rs.num_leaves != spec.num_leaves:
            raise ValueError("tree_map: tree structures do not match")
        cols.append(rl)
    return spec.unflatten([func(*xs) for xs in zip(*cols)])


# Exported as method-wrappers (opaque to Thunder's interpreter).
register_pytree_node = _register_pytree_node.__call__
tree_flatten = _tree_flatten.__call__
tree_unflatten = _tree_unflatten.__call__
tree_iter = _tree_iter.__call__
tree_leaves = _tree_leaves.__call__
tree_structure = _tree_structure.__call__
tree_map = _tree_map.__call__
