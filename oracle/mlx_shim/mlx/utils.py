"""mlx.utils on the NumPy shim: the one helper the reference's codec files import at module level."""


def tree_flatten(tree, prefix="", is_leaf=None):
    """Flatten nested dicts / lists / tuples into [(dotted.key, leaf), ...] (mlx.utils.tree_flatten)."""
    out = []
    if isinstance(tree, dict):
        items = tree.items()
    elif isinstance(tree, (list, tuple)):
        items = ((str(i), v) for i, v in enumerate(tree))
    else:
        return [(prefix, tree)]
    for k, v in items:
        key = f"{prefix}.{k}" if prefix else str(k)
        out.extend(tree_flatten(v, key, is_leaf))
    return out
