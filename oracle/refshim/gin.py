"""Minimal stand-in for gin-config, used ONLY to import the unmodified reference
(`/root/reference/featurePnP/model.py`) in the build container where gin-config is not
installed.  Test infrastructure: the product never imports this file.

`gin.configurable` / `gin.register` are used by the reference both bare (`@gin.configurable`)
and, elsewhere in the reference, in called form (`@gin.configurable("name")`); both become
pass-through decorators here (no argument injection).
"""


def _passthrough(*dargs, **dkwargs):
    if len(dargs) == 1 and callable(dargs[0]) and not dkwargs:
        return dargs[0]

    def deco(fn):
        return fn

    return deco


configurable = _passthrough
register = _passthrough
external_configurable = _passthrough
REQUIRED = object()


def parse_config_file(*a, **k):
    raise RuntimeError("gin stub: config parsing is not available")


def operative_config_str():
    return ""
