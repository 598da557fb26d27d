"""Registry seam (reference registry.py:17): when the reference tree (and mmengine) is importable
the modules register into ITS ``MODELS`` so ``MODELS.build(cfg.model)`` picks them up unchanged;
otherwise a dict-backed stand-in with the same two calls is used."""
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


try:  # pragma: no cover - only inside the reference's environment
    from registry import MODELS as _REF_MODELS  # type: ignore
    MODELS = _REF_MODELS
    USING_REFERENCE_REGISTRY = True
except Exception:
    MODELS = _LocalRegistry("model")
    USING_REFERENCE_REGISTRY = False


def register(cls):
    """Register under the reference's class name, replacing the reference class if present."""
    try:
        return MODELS.register_module(force=True)(cls)
    except TypeError:
        return MODELS.register_module()(cls)
