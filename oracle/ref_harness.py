"""Test infrastructure ONLY: imports the UNMODIFIED reference model from /root/reference.

The reference (`vilbert_k3m/vilbert_k3m.py`) imports `.utils`, which needs boto3 / botocore /
tensorboardX / torch._six at import time (`vilbert_k3m/utils.py:19-28`).  None of them is
installed, none is used on the pretraining path, so we pre-seed `sys.modules` with empty stubs.
Nothing in the reference tree is edited or copied.

This module only works where /root/reference exists (the build container).  The GPU box has no
reference tree: everything that runs there uses `oracle/k3m_oracle.py` (the restatement) and
the committed fixtures under `tests/golden/`.
"""
import math
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("K3M_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "vilbert_k3m", "vilbert_k3m.py"))


def _stub(name, **attrs):
    if name in sys.modules:
        return
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m


def import_reference():
    """Returns the reference module `vilbert_k3m.vilbert_k3m` (unmodified)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _stub("boto3")
    _stub("botocore")
    _stub("botocore.exceptions", ClientError=Exception)
    _stub("tensorboardX", SummaryWriter=object)
    _stub("torch._six", inf=math.inf, string_classes=(str,))
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib
    return importlib.import_module("vilbert_k3m.vilbert_k3m")


def make_reference_model(config_dict, driver_overrides=None):
    """Builds `BertForMultiModalPreTraining_tri_stru` the way `pretrain.py:1349-1366` does:
    BertConfig.from_dict(json) then attribute patches from the driver flags."""
    ref = import_reference()
    cfg = ref.BertConfig.from_dict(dict(config_dict))
    for k, v in (driver_overrides or {}).items():
        setattr(cfg, k, v)
    model = ref.BertForMultiModalPreTraining_tri_stru(cfg)
    return ref, cfg, model
