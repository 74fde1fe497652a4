"""Config helper mirroring lib/stnls/utils/config.py:5-24 (extract_pairs) without easydict."""
import copy


class AttrDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def extract_pairs(_cfg, defaults, restrict=True):
    """restrict=True : only the keys of `defaults` that are present in `_cfg` are extracted
                       (defaults are NOT filled in);
       restrict=False: `_cfg` is copied and missing keys are filled with the defaults.
    Same semantics as the reference (lib/stnls/utils/config.py:5-24)."""
    cfg = {} if restrict else copy.deepcopy(dict(_cfg))
    for key in defaults:
        if key in _cfg:
            cfg[key] = _cfg[key]
        elif not restrict:
            cfg[key] = defaults[key]
    return AttrDict(cfg)
