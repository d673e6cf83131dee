"""CPU: the compare-exchange programs of the separable 7 x 7 median (b2d_median7.cu) are generated code.  The generator
proves every merge / selection program by the 0-1 principle over all sorted inputs and the whole scheme on random
neighbourhoods (ties included) against the plain median; this test runs those proofs and checks that the committed
b2d_median7_nets.inc is what the generator writes."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _generator():
    spec = importlib.util.spec_from_file_location("gen_median7_networks", os.path.join(ROOT, "scripts", "gen_median7_networks.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_networks_are_proven_and_the_include_is_current(capsys):
    g = _generator()
    argv, sys.argv = sys.argv, ["gen_median7_networks.py"]
    try:
        nets = g.main()                      # asserts inside: 0-1 proofs + 300 random 8 x 8 neighbourhoods
    finally:
        sys.argv = argv
    out = capsys.readouterr().out
    assert "scheme verified on 300 random 8 x 8" in out and "5 x 5 scheme verified" in out
    with open(g.INC_PATH) as f:
        assert f.read() == g.render(nets), "b2d_median7_nets.inc is stale: python scripts/gen_median7_networks.py --write"
    ops = {k: len(v[0].ops) for k, v in nets.items()}
    assert ops == {"SORT6": 24, "MERGE33": 12, "PAIR": 34, "QUAD": 90, "CORE": 92, "ROW": 46, "MID2": 18,
                   "SORT4": 10, "PAIR5": 18, "CORE5": 40, "ROW5": 30, "MID5": 14}, ops
