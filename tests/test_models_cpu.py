"""CPU models of the RL kernels' per-tile arithmetic (tools/model_rle.py: rows, warp- and tile-level closed-form 255-splits,
staged positions; tools/model_rld.py: start-bit mask + PRMT selector table of the short-run expand) against the oracle /
numpy.  They are what the CUDA code was written from; no GPU needed."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "tools", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_rl_encode_tile_model_matches_oracle(capsys):
    _load("model_rle").main()
    assert "model ok" in capsys.readouterr().out


def test_rl_decode_short_run_model_matches_numpy(capsys):
    _load("model_rld").main()
    assert "model ok" in capsys.readouterr().out
