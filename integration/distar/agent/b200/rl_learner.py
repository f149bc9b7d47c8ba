"""learner (reference: distar/agent/default/rl_learner.py:60-71 constructs the model and the loss; `_train`, :82-135,
calls rl_learner_forward / compute_loss / backward / sync_gradients / its own GradClip and Adam unchanged)."""
from distar.agent.default import rl_learner as _ref
from applestar_b200 import Model, ReinforcementLoss

_ref.Model, _ref.ReinforcementLoss = Model, ReinforcementLoss
RLLearner = _ref.RLLearner
