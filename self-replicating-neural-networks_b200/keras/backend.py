"""`keras.backend` shim: the setups call clear_session() after every trial to drop the TF graph
(code/setups/network_trajectorys.py:27, mixed-soup.py:91, ...).  Here it is a no-op: there is no graph, and arena rows are recycled by reference counting."""


def clear_session():
    """Nothing to clear: arena rows are recycled as soon as their net object dies."""
    return None
