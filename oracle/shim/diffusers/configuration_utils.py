"""ConfigMixin / register_to_config: record constructor kwargs (diffusers semantics)."""
import functools
import inspect
import json


class FrozenDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class ConfigMixin:
    config_name = "config.json"

    def register_to_config(self, **kwargs):
        cfg = dict(getattr(self, "_internal_dict", {}))
        cfg.update(kwargs)
        self._internal_dict = FrozenDict(cfg)

    @property
    def config(self):
        return self._internal_dict

    @classmethod
    def from_config(cls, config, **kwargs):
        cfg = {k: v for k, v in dict(config).items() if not k.startswith("_")}
        cfg.update(kwargs)
        sig = inspect.signature(cls.__init__).parameters
        has_var_kw = any(p.kind == p.VAR_KEYWORD for p in sig.values())
        if not has_var_kw:
            cfg = {k: v for k, v in cfg.items() if k in sig}
        return cls(**cfg)

    @classmethod
    def load_config(cls, path, **kwargs):
        import os
        p = path if path.endswith(".json") else os.path.join(path, cls.config_name)
        with open(p) as f:
            return json.load(f)

    def save_config(self, save_directory, **kwargs):
        import os
        os.makedirs(save_directory, exist_ok=True)
        with open(os.path.join(save_directory, self.config_name), "w") as f:
            json.dump({"_class_name": type(self).__name__, **dict(self.config)}, f, indent=2, default=str)


def register_to_config(init):
    @functools.wraps(init)
    def inner(self, *args, **kwargs):
        sig = inspect.signature(init)
        params = [p for n, p in sig.parameters.items() if n != "self"]
        cfg = {}
        for p, a in zip(params, args):
            cfg[p.name] = a
        for p in params:
            if p.kind in (p.VAR_KEYWORD, p.VAR_POSITIONAL):
                continue
            if p.name not in cfg:
                cfg[p.name] = kwargs.get(p.name, p.default)
        init(self, *args, **kwargs)
        self.register_to_config(**cfg)
    return inner
