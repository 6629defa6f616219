"""Global config dict, mirroring the reference's ``utils.hparams.hparams`` (utils/hparams.py:13).

The hot path reads keys from it at construct- and call-time exactly where the reference does
(SURVEY.md section 5).  When this package is used INSIDE a DiffSinger checkout the
reference's own dict is adopted so that ``scripts/infer.py`` mutations are seen here.
"""
import sys

hparams = {}


def get_hparams() -> dict:
    mod = sys.modules.get("utils.hparams")
    ref = getattr(mod, "hparams", None) if mod is not None else None
    return ref if isinstance(ref, dict) and (ref or not hparams) else hparams
