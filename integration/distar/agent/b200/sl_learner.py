"""supervised learner (reference: distar/agent/default/sl_learner.py:14-16,37-38 construct the model and the loss; `_train`,
:46-53, calls model.sl_train(**data, hidden_state=...) and loss.compute_loss(logits, action_info, action_mask,
selected_units_num, entity_num, infer_action_info) unchanged)."""
from distar.agent.default import sl_learner as _ref
from applestar_b200 import Model, SupervisedLoss

_ref.Model, _ref.SupervisedLoss = Model, SupervisedLoss
SLLearner = _ref.SLLearner
