def maybe_num_nodes(index, num_nodes=None):
    """SURVEY A.2.  ``num_nodes=None`` costs a device->host sync, as in the reference."""
    return int(index.max().item()) + 1 if num_nodes is None else num_nodes
