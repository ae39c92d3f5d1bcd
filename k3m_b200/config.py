"""Model configuration for the K3M pretraining step.

Mirrors the reference's `BertConfig` surface (`vilbert_k3m/vilbert_k3m.py:149-308`): same
constructor keywords and defaults, `from_dict` / `from_json_file` / `to_dict` /
`to_json_string` / `__repr__`, and the same quirk that `from_dict` first builds a config with
every constructor default and then lets JSON keys override attributes one by one
(`vilbert_k3m.py:284-289`).  Attributes the reference's drivers inject after loading the JSON
(`pretrain.py:1350-1366`) get the drivers' defaults here so that a bare JSON is enough.
"""
import copy
import json

# constructor defaults, `vilbert_k3m.py:153-196`
_CTOR_DEFAULTS = dict(
    hidden_size=768,
    num_hidden_layers=12,
    num_attention_heads=12,
    intermediate_size=3072,
    hidden_act="gelu",
    hidden_dropout_prob=0.1,
    attention_probs_dropout_prob=0.1,
    max_position_embeddings=512,
    type_vocab_size=2,
    initializer_range=0.02,
    v_feature_size=2048,
    v_target_size=1601,
    v_hidden_size=768,
    v_num_hidden_layers=3,
    v_num_attention_heads=12,
    v_intermediate_size=3072,
    bi_hidden_size=1024,
    bi_num_attention_heads=16,
    v_attention_probs_dropout_prob=0.1,
    v_hidden_act="gelu",
    v_hidden_dropout_prob=0.1,
    v_initializer_range=0.2,
    v_biattention_id=[0, 1],
    t_biattention_id=[10, 11],
    visual_target=0,
    fast_mode=False,
    fixed_v_layer=0,
    fixed_t_layer=0,
    in_batch_pairs=False,
    fusion_method="mul",
    dynamic_attention=False,
    with_coattention=True,
    objective=0,
    num_negative_image=128,
    num_negative_pv=4,
    margin=1.0,
    model="bert",
    task_specific_tokens=False,
    visualization=False,
    use_image=True,
)

# attributes that only the drivers set (`pretrain.py:1350-1366`, flag defaults `:1262-1328`)
_DRIVER_DEFAULTS = dict(if_pre_sampling=1, loss_type="ce")


class BertConfig(object):
    """Configuration of the K3M model; drop-in for the reference class of the same name."""

    def __init__(self, vocab_size_or_config_json_file, **kwargs):
        unknown = set(kwargs) - set(_CTOR_DEFAULTS)
        if unknown:
            raise TypeError("unexpected BertConfig arguments: %s" % sorted(unknown))
        args = dict(_CTOR_DEFAULTS)
        args.update(kwargs)
        # same three asserts as the reference, `vilbert_k3m.py:222-224`
        assert len(args["v_biattention_id"]) == len(args["t_biattention_id"])
        assert max(args["v_biattention_id"]) < args["v_num_hidden_layers"]
        assert max(args["t_biattention_id"]) < args["num_hidden_layers"]
        if isinstance(vocab_size_or_config_json_file, str):
            with open(vocab_size_or_config_json_file, "r", encoding="utf-8") as reader:
                for key, value in json.loads(reader.read()).items():
                    self.__dict__[key] = value
        elif isinstance(vocab_size_or_config_json_file, int):
            self.vocab_size = vocab_size_or_config_json_file
            for key, value in args.items():
                self.__dict__[key] = copy.deepcopy(value)
        else:
            raise ValueError(
                "First argument must be either a vocabulary size (int)"
                "or the path to a pretrained model config file (str)"
            )
        for key, value in _DRIVER_DEFAULTS.items():
            self.__dict__.setdefault(key, value)

    @classmethod
    def from_dict(cls, json_object):
        config = BertConfig(vocab_size_or_config_json_file=-1)
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-8") as reader:
            text = reader.read()
        return cls.from_dict(json.loads(text))

    def __repr__(self):
        return str(self.to_json_string())

    def to_dict(self):
        return copy.deepcopy(self.__dict__)

    def to_json_string(self):
        return json.dumps(self.to_dict(), indent=2, sort_keys=True) + "\n"


# The only config the reference ships (`config/bert_base_6layer_6conect.json`), restated as a
# dict so that the GPU box (no reference tree) can build the headline model.
K3M_BASE_6LAYER_6CONNECT = {
    "attention_probs_dropout_prob": 0.1,
    "hidden_act": "gelu",
    "hidden_dropout_prob": 0.1,
    "hidden_size": 768,
    "initializer_range": 0.02,
    "intermediate_size": 3072,
    "max_position_embeddings": 512,
    "num_attention_heads": 12,
    "num_hidden_layers": 12,
    "type_vocab_size": 2,
    "vocab_size": 21128,
    "v_feature_size": 2048,
    "v_target_size": 1601,
    "v_hidden_size": 1024,
    "v_num_hidden_layers": 6,
    "v_num_attention_heads": 8,
    "v_intermediate_size": 1024,
    "bi_hidden_size": 1024,
    "bi_num_attention_heads": 8,
    "bi_intermediate_size": 1024,
    "bi_attention_type": 1,
    "v_attention_probs_dropout_prob": 0.1,
    "v_hidden_act": "gelu",
    "v_hidden_dropout_prob": 0.1,
    "v_initializer_range": 0.02,
    "v_biattention_id": [0, 1, 2, 3, 4, 5],
    "t_biattention_id": [6, 7, 8, 9, 10, 11],
    "pooling_method": "mul",
}

# A small config with the same topology (two connection layers) for fast parity tests.  All
# widths are multiples of 64 so the sm_100a GEMM tiles apply unchanged.
K3M_TINY = {
    "attention_probs_dropout_prob": 0.1,
    "hidden_act": "gelu",
    "hidden_dropout_prob": 0.1,
    "hidden_size": 128,
    "initializer_range": 0.02,
    "intermediate_size": 256,
    "max_position_embeddings": 64,
    "num_attention_heads": 2,
    "num_hidden_layers": 4,
    "type_vocab_size": 2,
    "vocab_size": 320,
    "v_feature_size": 64,
    "v_target_size": 41,
    "v_hidden_size": 256,
    "v_num_hidden_layers": 3,
    "v_num_attention_heads": 2,
    "v_intermediate_size": 256,
    "bi_hidden_size": 256,
    "bi_num_attention_heads": 2,
    "v_attention_probs_dropout_prob": 0.1,
    "v_hidden_act": "gelu",
    "v_hidden_dropout_prob": 0.1,
    "v_initializer_range": 0.02,
    "v_biattention_id": [0, 1],
    "t_biattention_id": [2, 3],
}
