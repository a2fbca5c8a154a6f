"""CPU suite: the oracle reproduces tests/golden/digests_1m.json -- the per-step digests that
tests/golden/make_digests_1m.py wrote while the REFERENCE (imported in place from /root/reference) was stepping
>= 1,048,576 seeded trajectories per integer env head to head with the oracle.  The GPU suite asserts the same file
from the CUDA kernels (tests/test_gpu_million.py), so reference -> oracle -> CUDA are chained on the same trajectories.

`PULSE_REFSLOW=1 pytest -m refslow` re-runs the head-to-head against the live reference (build container only)."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from tests import digest_util as du


class OracleEngine:
    def __init__(self, hr, name):
        c = du.CASES[name]
        self.c, self.n = c, c["runs"] * c["n_run"]
        self.orc = pyoracle.OraclePoker(hr, n_players=c["P"], max_players=c["M"], n_games=self.n, **c["cfg"])

    def reset(self, hand, ap, rot):
        self.orc.tape_decks(self.c["seed"], 0, hand)
        return self.orc.reset(None, active_players=ap, rotation=rot)

    def step(self, a):
        return self.orc.step(a)

    def get(self, f):
        return getattr(self.orc, f)


@pytest.mark.parametrize("name", ["poker_hu", "poker_6max", "poker_10max"])
def test_oracle_reproduces_reference_digests_poker(name, hr_table):
    assert du.check_poker_case(name, OracleEngine(hr_table, name)) >= 1 << 20


def test_oracle_reproduces_reference_digests_blackjack():
    c, want = du.CASES["blackjack"], du.load_digests()["blackjack"]["episode"]
    n = c["runs"] * c["n_run"]
    orc = pyoracle.OracleBlackjack(n)
    orc.tape_decks(c["seed"], 0, 0)
    assert du.hexd(du.digest(orc.reset(None))) == want["reset_obs"]
    acts = du.case_actions("blackjack")
    for t in range(c["steps"]):
        o, r, d = orc.step(acts[t])
        w = want["steps"][t]
        assert du.hexd(du.digest(o)) == w["obs"] and du.hexd(du.digest(r)) == w["rewards"], f"blackjack step {t}"
        assert du.hexd(du.digest(np.asarray(d).astype(np.uint8))) == w["terminated"] and int(d.sum()) == w["terminated_count"]


def test_oracle_reproduces_reference_digests_tfe():
    c, want = du.CASES["tfe_4x4"], du.load_digests()["tfe_4x4"]["episode"]
    n, seed = c["n"], c["seed"]
    orc = pyoracle.OracleTFE(n, c["side"])
    assert du.hexd(du.digest(orc.reset(orc.tape(seed, 0, pyoracle.STREAM_TFE_RESET, 0, 4)))) == want["reset_boards"]
    acts = du.case_actions("tfe_4x4")
    for t in range(c["steps"]):
        b, r, d = orc.step(acts[t], orc.tape(seed, 0, pyoracle.STREAM_TFE_SPAWN, t + 1, 2))
        w = want["steps"][t]
        assert du.hexd(du.digest(b)) == w["boards"] and du.hexd(du.digest(r)) == w["rewards"], f"tfe step {t}"
        assert du.hexd(du.digest(d.astype(np.uint8))) == w["done"] and du.hexd(du.digest(orc.scores)) == w["scores"]


@pytest.mark.refslow
@pytest.mark.skipif(os.environ.get("PULSE_REFSLOW") != "1", reason="set PULSE_REFSLOW=1 (build container, ~25 min)")
def test_reference_head_to_head_regenerates_the_same_file(tmp_path):
    """Runs tests/golden/make_digests_1m.py against the live reference and checks it writes the committed digests."""
    import json
    import subprocess
    import sys
    from pathlib import Path

    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("/root/reference not present")
    root = Path(__file__).resolve().parents[1]
    golden = root / "tests" / "golden" / "digests_1m.json"
    before = json.loads(golden.read_text())
    subprocess.run([sys.executable, str(root / "tests" / "golden" / "make_digests_1m.py")], check=True, cwd=root)
    after = json.loads(golden.read_text())
    for name in before:
        for k in ("hands", "episode"):
            if k in before[name]:
                assert before[name][k] == after[name][k], name
