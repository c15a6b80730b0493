"""ruamel.yaml stand-in over PyYAML (safe_load / dump / YAML().load)."""
import yaml as _y

safe_load = _y.safe_load
safe_dump = _y.safe_dump
YAMLError = _y.YAMLError


def load(stream, Loader=None):
    return _y.safe_load(stream)


def dump(data, stream=None, **kw):
    kw.pop("Dumper", None)
    return _y.safe_dump(data, stream, **kw)


class YAML:
    def __init__(self, typ=None, pure=False):
        pass

    def load(self, stream):
        return _y.safe_load(stream)

    def dump(self, data, stream=None):
        return _y.safe_dump(data, stream)
