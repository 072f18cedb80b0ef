"""Pass-through stand-ins for gin.configurable / gin.register, used only when gin-config is not
installed.  With real gin present (`import gin` succeeds) this module is never imported, and the
reference's `.gin` files bind `sparseFeaturePnP`, `find_inliers` and the losses as before."""


def _passthrough(*dargs, **dkwargs):
    if len(dargs) == 1 and callable(dargs[0]) and not dkwargs:
        return dargs[0]

    def deco(fn):
        return fn

    return deco


configurable = _passthrough
register = _passthrough
external_configurable = _passthrough
