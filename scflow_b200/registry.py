"""Registry seam (reference registry.py:17).  The package keeps its own dict-backed registry with the same two calls
(`register_module`, `build`); `install()` is the explicit, opt-in step that overrides the reference's classes in the
reference's own ``MODELS``."""
from __future__ import annotations


class _LocalRegistry:
    def __init__(self, name):
        self.name = name
        self._modules = {}

    def register_module(self, name=None, force=False, module=None):
        def deco(cls):
            key = name or cls.__name__
            if key in self._modules and not force:
                raise KeyError(f"{key} is already registered in {self.name}")
            self._modules[key] = cls
            return cls
        return deco if module is None else deco(module)

    def get(self, key):
        return self._modules.get(key)

    def build(self, cfg):
        cfg = dict(cfg)
        t = cfg.pop("type")
        cls = self._modules[t] if isinstance(t, str) else t
        return cls(**cfg)


MODELS = _LocalRegistry("model")
_CLASSES = []


def register(cls):
    """Register into scflow_b200's own registry (used by `MODELS.build(pose_head_cfg)` inside the decoder).
    Nothing outside this package is touched at import time; see `install`."""
    _CLASSES.append(cls)
    return MODELS.register_module(force=True)(cls)


def install(models_registry=None):
    """Opt-in override inside the reference's environment: put SCFlowDecoder / MultiClassPoseHead / RAFTEncoder
    into the reference's `registry.MODELS` (registry.py:17) under the reference's class names (force=True), so
    `MODELS.build(cfg.model)` builds these instead.  Call it AFTER the reference's `models` package has been
    imported and only for configs on the supported path (net_type='Basic', 6-D rotation, mask_corr=False ...):
    the replaced classes raise on anything else, which is why importing scflow_b200 alone does not do this.
    Returns the names it replaced."""
    if models_registry is None:
        from registry import MODELS as models_registry  # type: ignore  # the reference tree must be importable
    for cls in _CLASSES:
        try:
            models_registry.register_module(force=True, module=cls)
        except TypeError:                       # a registry without `module=`: decorator form
            models_registry.register_module(force=True)(cls)
    return [c.__name__ for c in _CLASSES]
