"""CPU test (skipped where the reference checkout is absent, e.g. on the GPU box): every name the reference's setup
scripts use after their star-imports resolves against the drop-in modules -- the static half of "code/setups/*.py run
unchanged" (the dynamic half, the same flows on a GPU, is tests/test_gpu_setups.py)."""
import ast
import builtins
import glob
import os

import pytest

REF = "/root/reference/code/setups"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")


def _free_names(tree):
    """Names loaded at any level that are not bound anywhere in the module (assignments, defs, imports, args, loops)."""
    bound, loaded = set(), set()
    for node in ast.walk(tree):
        if isinstance(node, ast.Name):
            (loaded if isinstance(node.ctx, ast.Load) else bound).add(node.id)
        elif isinstance(node, (ast.FunctionDef, ast.ClassDef)):
            bound.add(node.name)
        elif isinstance(node, ast.arg):
            bound.add(node.arg)
        elif isinstance(node, (ast.Import, ast.ImportFrom)):
            for alias in node.names:
                if alias.name != "*":
                    bound.add((alias.asname or alias.name).split(".")[0])
        elif isinstance(node, ast.ExceptHandler) and node.name:
            bound.add(node.name)
    return loaded - bound - set(dir(builtins))


@pytest.mark.parametrize("script", sorted(glob.glob(os.path.join(REF, "*.py"))))
def test_setup_script_names_resolve(script):
    import experiment
    import network
    import soup
    import util
    tree = ast.parse(open(script).read())
    star_from = {node.module for node in ast.walk(tree) if isinstance(node, ast.ImportFrom)
                 and any(a.name == "*" for a in node.names)}
    assert star_from <= {"util", "experiment", "network", "soup"}
    namespace = set()
    for mod, name in ((util, "util"), (experiment, "experiment"), (network, "network"), (soup, "soup")):
        if name in star_from:
            public = getattr(mod, "__all__", None) or [n for n in vars(mod) if not n.startswith("_")]
            namespace |= set(public)
    missing = _free_names(tree) - namespace
    assert not missing, "%s uses names the drop-in does not provide: %s" % (os.path.basename(script), sorted(missing))
    # explicit imports from the drop-in modules
    for node in ast.walk(tree):
        if isinstance(node, ast.ImportFrom) and node.module in ("soup", "network", "experiment", "util"):
            for alias in node.names:
                if alias.name != "*":
                    assert hasattr({"soup": soup, "network": network, "experiment": experiment, "util": util}[node.module],
                                   alias.name), (script, alias.name)
    import keras.backend
    assert callable(keras.backend.clear_session)
