"""`keras.backend` shim: the setups call clear_session() after every trial to drop the TF graph
(code/setups/network_trajectorys.py:27, mixed-soup.py:91, ...).  Here it only lets the arena recycle rows."""
import gc


def clear_session():
    gc.collect()
