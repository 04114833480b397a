"""flax.struct.dataclass: a frozen dataclass with .replace()."""
import dataclasses


def dataclass(cls):
    cls = dataclasses.dataclass(frozen=True)(cls)

    def replace(self, **kw):
        return dataclasses.replace(self, **kw)

    cls.replace = replace
    return cls
