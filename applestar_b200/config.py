"""Model / loss configuration of the hot path, as plain Python.

Mirrors the keys callers read from the reference's merged EasyDict config
(distar/agent/default/model/actor_critic_default_config.yaml,
distar/agent/default/rl_training/default_reinforcement_loss.yaml and the learner block of
distar/bin/rl_user_config.yaml:41,58-117). Only what the policy/value forward-backward path
needs is kept; field order matters because it fixes channel order of the concatenated
embeddings.
"""
import copy

SPATIAL_Y, SPATIAL_X = 152, 160          # lib/features.py:31
MAX_ENTITY_NUM = 512                      # lib/features.py:37
MAX_SELECTED_UNITS_NUM = 64               # lib/features.py:36
EFFECT_LEN = 100                          # lib/features.py:38
NUM_ACTIONS = 327
NUM_UNIT_TYPES = 260
NUM_UPGRADES = 90
NUM_QUEUE_ACTIONS = 49
NUM_BEGINNING_ORDER_ACTIONS = 174
NUM_CUMULATIVE_STAT_ACTIONS = 167
NUM_UNIT_MIX_ABILITIES = 269
NUM_BUFFS = 50
NUM_ADDON = 9
BEGINNING_ORDER_LENGTH = 20

# (name, storage dtype, arc, num_embeddings) in yaml order  -- 997 columns in total
# dtype per lib/features.py:57-67
ENTITY_FIELDS = [
    ('unit_type', 'int16', 'one_hot', NUM_UNIT_TYPES),
    ('alliance', 'uint8', 'one_hot', 5),
    ('cargo_space_taken', 'uint8', 'one_hot', 9),
    ('build_progress', 'float16', 'unsqueeze', 1),
    ('health_ratio', 'float16', 'unsqueeze', 1),
    ('shield_ratio', 'float16', 'unsqueeze', 1),
    ('energy_ratio', 'float16', 'unsqueeze', 1),
    ('display_type', 'uint8', 'one_hot', 5),
    ('x', 'uint8', 'binary', 11),
    ('y', 'uint8', 'binary', 11),
    ('cloak', 'uint8', 'one_hot', 5),
    ('is_blip', 'uint8', 'one_hot', 2),
    ('is_powered', 'uint8', 'one_hot', 2),
    ('mineral_contents', 'float16', 'unsqueeze', 1),
    ('vespene_contents', 'float16', 'unsqueeze', 1),
    ('cargo_space_max', 'uint8', 'one_hot', 9),
    ('assigned_harvesters', 'uint8', 'one_hot', 24),
    ('weapon_cooldown', 'uint8', 'one_hot', 32),
    ('order_length', 'uint8', 'one_hot', 9),
    ('order_id_0', 'int16', 'one_hot', NUM_ACTIONS),
    ('order_id_1', 'int16', 'one_hot', NUM_QUEUE_ACTIONS),
    ('is_hallucination', 'uint8', 'one_hot', 2),
    ('buff_id_0', 'uint8', 'one_hot', NUM_BUFFS),
    ('buff_id_1', 'uint8', 'one_hot', NUM_BUFFS),
    ('addon_unit_type', 'uint8', 'one_hot', NUM_ADDON),
    ('is_active', 'uint8', 'one_hot', 2),
    ('order_progress_0', 'float16', 'unsqueeze', 1),
    ('order_progress_1', 'float16', 'unsqueeze', 1),
    ('order_id_2', 'int16', 'one_hot', NUM_QUEUE_ACTIONS),
    ('order_id_3', 'int16', 'one_hot', NUM_QUEUE_ACTIONS),
    ('is_in_cargo', 'uint8', 'one_hot', 2),
    ('attack_upgrade_level', 'uint8', 'one_hot', 4),
    ('armor_upgrade_level', 'uint8', 'one_hot', 4),
    ('shield_upgrade_level', 'uint8', 'one_hot', 4),
    ('last_selected_units', 'int8', 'one_hot', 2),
    ('last_targeted_unit', 'int8', 'one_hot', 2),
]
ENTITY_INPUT_DIM = sum(n for _, _, _, n in ENTITY_FIELDS)
assert ENTITY_INPUT_DIM == 997

# (name, arc, num_embeddings) in yaml order; effects are 100-slot index lists (int16)
SPATIAL_FIELDS = [
    ('height_map', 'other', 1),
    ('visibility_map', 'one_hot', 4),
    ('creep', 'one_hot', 2),
    ('player_relative', 'one_hot', 5),
    ('alerts', 'one_hot', 2),
    ('pathable', 'one_hot', 2),
    ('buildable', 'one_hot', 2),
    ('effect_PsiStorm', 'scatter', 1),
    ('effect_NukeDot', 'scatter', 1),
    ('effect_LiberatorDefenderZone', 'scatter', 1),
    ('effect_BlindingCloud', 'scatter', 1),
    ('effect_CorrosiveBile', 'scatter', 1),
    ('effect_LurkerSpines', 'scatter', 1),
]

# (name, arc, in/num_embeddings, out, scalar_context, baseline_feature) in yaml order
SCALAR_FIELDS = [
    ('agent_statistics', 'fc', 10, 64, False, True),
    ('home_race', 'one_hot', 5, 32, True, False),
    ('away_race', 'one_hot', 5, 32, True, False),
    ('upgrades', 'fc', NUM_UPGRADES, 128, False, True),
    ('time', 'identity', 0, 32, False, False),
    ('unit_counts_bow', 'fc', NUM_UNIT_TYPES, 128, False, True),
    ('last_delay', 'one_hot', 128, 64, False, False),
    ('last_queued', 'one_hot', 2, 32, False, False),
    ('last_action_type', 'one_hot', NUM_ACTIONS, 128, False, False),
    ('cumulative_stat', 'fc', NUM_CUMULATIVE_STAT_ACTIONS, 128, True, True),
    ('beginning_order', 'transformer', NUM_BEGINNING_ORDER_ACTIONS, 64, True, True),
    ('unit_type_bool', 'fc', NUM_UNIT_TYPES, 64, True, False),
    ('enemy_unit_type_bool', 'fc', NUM_UNIT_TYPES, 64, True, False),
    ('unit_order_type', 'fc', NUM_UNIT_MIX_ABILITIES, 64, True, False),
]
# storage dtypes (lib/features.py:44-54)
SCALAR_DTYPES = {
    'home_race': ('uint8', ()), 'away_race': ('uint8', ()), 'upgrades': ('int16', (NUM_UPGRADES,)),
    'time': ('float32', ()), 'unit_counts_bow': ('uint8', (NUM_UNIT_TYPES,)),
    'agent_statistics': ('float32', (10,)), 'cumulative_stat': ('uint8', (NUM_CUMULATIVE_STAT_ACTIONS,)),
    'beginning_order': ('int16', (BEGINNING_ORDER_LENGTH,)), 'last_queued': ('int16', ()),
    'last_delay': ('int16', ()), 'last_action_type': ('int16', ()),
    'bo_location': ('int16', (BEGINNING_ORDER_LENGTH,)),
    'unit_order_type': ('uint8', (NUM_UNIT_MIX_ABILITIES,)), 'unit_type_bool': ('uint8', (NUM_UNIT_TYPES,)),
    'enemy_unit_type_bool': ('uint8', (NUM_UNIT_TYPES,)),
}

VALUE_FIELDS = ['winloss', 'build_order', 'built_unit', 'effect', 'upgrade', 'battle']
HEADS = ['action_type', 'delay', 'queued', 'selected_units', 'target_unit', 'target_location']


class AttrDict(dict):
    """dict with attribute access, recursive (what callers expect of the reference's EasyDict)."""

    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = v

    @staticmethod
    def _wrap(v):
        if isinstance(v, dict) and not isinstance(v, AttrDict):
            return AttrDict(v)
        return v

    def __setitem__(self, k, v):
        super().__setitem__(k, self._wrap(v))

    def __setattr__(self, k, v):
        self[k] = v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __deepcopy__(self, memo):
        return AttrDict({k: copy.deepcopy(v, memo) for k, v in self.items()})


def deep_merge(base, override):
    """Recursive dict merge, override wins (ctools/utils/config_helper.py:10-57 semantics)."""
    out = copy.deepcopy(dict(base))
    for k, v in (override or {}).items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = deep_merge(out[k], v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def _value_param(name, atan):
    return {'name': name, 'param': {'name': name, 'input_dim': 384, 'activation': 'relu', 'norm_type': 'LN',
                                    'res_dim': 256, 'res_num': 16, 'atan': atan}}


DEFAULT_MODEL_CONFIG = {
    'common': {'type': 'rl'},
    'learner': {'use_value_feature': False},
    'model': {
        'spatial_x': SPATIAL_X, 'spatial_y': SPATIAL_Y, 'temperature': 1.0,
        'enable_baselines': list(VALUE_FIELDS),
        'entity_reduce_type': 'selected_units_num',
        'only_update_baseline': False,
        'value': {f: _value_param(f, f == 'winloss') for f in VALUE_FIELDS},
        'encoder': {
            'obs_encoder': {
                'scalar_encoder': {'activation': 'relu', 'output_dim': 1024},
                'spatial_encoder': {'input_dim': 56, 'resblock_num': 4, 'fc_dim': 256, 'project_dim': 32,
                                    'downsample_type': 'maxpool', 'down_channels': [64, 128, 128],
                                    'activation': 'relu', 'norm_type': 'none', 'head_type': 'fc'},
                'entity_encoder': {'input_dim': 997, 'head_dim': 128, 'hidden_dim': 1024, 'output_dim': 256,
                                   'head_num': 2, 'mlp_num': 2, 'layer_num': 3, 'dropout_ratio': 0,
                                   'activation': 'relu', 'ln_type': 'post'},
            },
            'scatter': {'input_dim': 256, 'output_dim': 32, 'scatter_type': 'add'},
            'core_lstm': {'lstm_type': 'normal', 'input_size': 1536, 'hidden_size': 384, 'num_layers': 3,
                          'dropout': 0.0},
        },
        'policy': {'head': {
            'action_type_head': {'input_dim': 384, 'res_dim': 256, 'res_num': 2, 'action_num': NUM_ACTIONS,
                                 'action_map_dim': 256, 'gate_dim': 1024, 'context_dim': 448,
                                 'activation': 'relu', 'norm_type': 'LN'},
            'delay_head': {'input_dim': 1024, 'decode_dim': 256, 'delay_dim': 128, 'delay_map_dim': 256},
            'queued_head': {'input_dim': 1024, 'decode_dim': 256, 'queued_dim': 2, 'queued_map_dim': 256},
            'selected_units_head': {'input_dim': 1024, 'entity_embedding_dim': 256, 'key_dim': 32,
                                    'func_dim': 256, 'hidden_dim': 32, 'num_layers': 1},
            'target_unit_head': {'input_dim': 1024, 'entity_embedding_dim': 256, 'key_dim': 32},
            'location_head': {'input_dim': 1024, 'upsample_type': 'bilinear', 'upsample_dims': [64, 32, 1],
                              'res_dim': 128, 'res_num': 4, 'film': False, 'gate': True, 'unet': False,
                              'reshape_channel': 4, 'map_skip_dim': 128},
        }},
    },
}

_HW1 = {'action_type': 1.0, 'delay': 1.0, 'queued': 1.0, 'selected_units': 1.0, 'target_unit': 1.0,
        'target_location': 1.0}
_ZW = {f: 0.0 for f in VALUE_FIELDS}

# default_reinforcement_loss.yaml merged with the learner block of bin/rl_user_config.yaml
# (the user block supplies `queued`, `dapo*`, and overrides kl / head weights).
DEFAULT_LOSS_CONFIG = {
    'loss_weights': {
        'baseline': dict(_ZW, winloss=10.0), 'pg': dict(_ZW, winloss=1.0), 'upgo': {'winloss': 1.0},
        'kl': 0.002, 'kl_su': 1, 'action_type_kl': 0.1, 'dapo': 0.1, 'entropy': 0.0001,
    },
    'pg_head_weights': dict(_HW1), 'upgo_head_weights': dict(_HW1), 'entropy_head_weights': dict(_HW1),
    'kl_head_weights': dict(_HW1), 'dapo_head_weights': dict(_HW1),
    'kl': {'action_type_kl_steps': 5200},
    'dapo': {'dapo_steps': 2400},
    'use_dapo': False,
    'gammas': {'baseline': dict({f: 1.0 for f in VALUE_FIELDS}, battle=0.997),
               'pg': dict({f: 1.0 for f in VALUE_FIELDS}, battle=0.997)},
}
