"""TEST INFRASTRUCTURE ONLY.  Import stub for `omegaconf` (absent from this image): the reference's
src/common/diffusion/config.py and src/common/config.py only need the names at import time; the golden generator builds
schedules / samplers through their constructors, never through a YAML config."""


class DictConfig(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class ListConfig(list):
    pass


class OmegaConf:
    @staticmethod
    def create(obj=None):
        return DictConfig(obj or {})

    @staticmethod
    def load(path):
        raise NotImplementedError("omegaconf stub: YAML loading is not part of the oracle")

    @staticmethod
    def register_new_resolver(*a, **k):
        pass

    @staticmethod
    def to_container(cfg, **k):
        return dict(cfg)

    @staticmethod
    def to_object(cfg):
        return dict(cfg)

    @staticmethod
    def set_readonly(*a, **k):
        pass
