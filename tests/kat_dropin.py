"""Adapter that replays the reference's recorded unit-test calls (tests/golden/poker_reftests.npz) through the
drop-in env class `pulselib_b200.envs.poker.PokerGPU` (packed engine + mirror on access) on any device: the B200 in
tests/test_gpu_reference_kat.py, the host-compiled kernel bodies (tests/hostemu/shim.py) in tests/test_reference_kat.py."""
from unittest import mock

import numpy as np
import torch

from tests import poker_util as pu


class DropinKat:
    def __init__(self, dev, ctor):
        from pulselib_b200.envs.poker import PokerGPU

        self.dev, self.ctor = dev, ctor
        self.env = PokerGPU(device=dev, agents=[], **ctor)
        self.N, self.P = ctor["n_games"], ctor["n_players"]

    def _t(self, a, dtype):
        return torch.from_numpy(np.ascontiguousarray(a)).to(self.dev).to(dtype).contiguous()

    def set_state(self, before, active_players, rec):
        env = self.env
        if rec.method == "reset" and rec.first:
            return                                  # untouched env: stacks is None, no button yet
        env.reset(options={"active_players": False})  # allocates what the reference's reset allocates
        if rec.method == "reset" and not rec.had_button:
            del env.button_pos                      # stacks were preset by the test, no reset had run (:121)
        env.active_players = active_players
        for f in pu.KAT_N + pu.KAT_NP + ("decks", "board", "hands"):
            if f in before:
                setattr(env, f, self._t(before[f], torch.int32))
        for f in ("is_done", "equity_dirty"):
            if f in before:
                setattr(env, f, self._t(before[f], torch.bool))
        if "equities" in before:                    # the kernel's row stride is active_players
            env.equities = self._t(before["equities"][:, :active_players], torch.float32)

    def get_state(self, name):
        v = getattr(self.env, name)
        return v.detach().cpu().numpy() if torch.is_tensor(v) else v

    def reset(self, rec):
        a, ap = rec.args, int(rec.after["active_players"])
        opts = {"prefixed_decks": torch.from_numpy(rec.after["decks"])}   # the deck the reference drew or was given
        if int(a["opt_rotation"]) >= 0:
            opts["rotation"] = int(a["opt_rotation"])
        if int(a["opt_q_agent_seat"]) >= 0:
            opts["q_agent_seat"] = int(a["opt_q_agent_seat"])
        if int(a["opt_active_players"]) == 1:       # the table size the reference's randint gave
            opts["active_players"] = True
            with mock.patch("torch.randint", lambda *aa, **kk: torch.tensor([ap])):
                obs, _ = self.env.reset(options=opts, rotation=int(a["arg_rotation"]))
        else:
            obs, _ = self.env.reset(options=opts, rotation=int(a["arg_rotation"]))
        return {"obs": obs.cpu().numpy()}

    def call(self, rec):
        env, m, a = self.env, rec.method, rec.args
        if m == "step":
            if "stub_equities" in a:                # the test replaced calculate_equities on the instance
                eq = self._t(a["stub_equities"][:, : env.active_players], torch.float32)
                dirty = self._t(a["stub_equity_dirty"], torch.bool)

                def fake():
                    env.equities.copy_(eq)
                    env.equity_dirty.copy_(dirty)

                env.calculate_equities = fake
            obs, rew, done, _, _ = env.step(torch.from_numpy(a["a0"]).to(self.dev))
            return {"obs": obs.cpu().numpy(), "reward": rew.cpu().numpy(), "done": done.cpu().numpy().astype(np.uint8)}
        if m == "get_obs":
            return {"value": env.get_obs().cpu().numpy()}
        if m == "execute_actions":
            env.execute_actions(torch.from_numpy(a["a0"]).to(self.dev))
            return {}
        if m in ("calculate_equities", "resolve_fold_winners", "resolve_terminated_games", "post_blinds"):
            getattr(env, m)()
            return {}
        if m == "poker_reward_gpu":
            r = env.poker_reward_gpu(torch.from_numpy(a["a0"]).to(self.dev), torch.from_numpy(a["a1"]).to(self.dev))
            return {"value": r.cpu().numpy()}
        if m == "deal_players_cards":
            return {"value": env.deal_players_cards(int(a["a0"])).cpu().numpy()}
        if m == "deal_cards":
            return {"value": env.deal_cards(torch.from_numpy(np.atleast_1d(a["a0"])).to(self.dev), int(a["a1"])).cpu().numpy()}
        raise AssertionError(f"unhandled reference method {m}")
