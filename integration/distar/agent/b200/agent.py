"""actors / evaluation (reference: distar/agent/default/agent.py:131-141 builds `Model(cfg)` and calls .cuda() / .eval())."""
from distar.agent.default import agent as _ref
from applestar_b200 import Model            # same constructor, compute_logp_action / compute_teacher_logit

_ref.Model = Model
Agent = _ref.Agent
