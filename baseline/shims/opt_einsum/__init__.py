"""Stand-in for ``opt_einsum``: the reference imports it at ``thunder/torch/__init__.py:18`` but only
``ltorch.einsum`` reaches it, which is off the fused-region hot path. Baseline infrastructure only."""


def contract_path(*args, **kwargs):
    raise NotImplementedError("opt_einsum is not available in this sandbox (einsum is off the hot path)")
