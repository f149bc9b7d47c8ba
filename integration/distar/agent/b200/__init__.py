"""Agent pipeline `b200` for the reference tree (copy this directory to distar/agent/b200, or put its parent on
`distar.agent.__path__`): the reference's own Agent / RLLearner with `Model` and `ReinforcementLoss` replaced by the
B200-native ones of `applestar_b200`.  Selected by the config keys `learner.agent: b200` / `actor.agents.<player>: b200`
through distar/agent/import_helper.py:3-14."""
