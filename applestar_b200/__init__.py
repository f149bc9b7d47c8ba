"""applestar_b200 -- B200-native (sm_100a) implementation of the DI-star / Applestar policy-value hot path.

Public surface mirrors the reference (`distar.agent.default.model.Model`,
`distar.agent.default.rl_training.rl_loss.ReinforcementLoss`, `distar.agent.default.sl_training.sl_loss.SupervisedLoss`); see INTEGRATION.md.
"""
from .model import Model  # noqa: F401
from .rl_loss import ReinforcementLoss  # noqa: F401
from .sl_loss import SupervisedLoss  # noqa: F401

__all__ = ['Model', 'ReinforcementLoss', 'SupervisedLoss']
