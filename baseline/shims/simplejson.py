"""simplejson stand-in over the stdlib json module."""
import json as _j
from json import JSONDecodeError, JSONEncoder, load, loads  # noqa: F401


def dumps(obj, *a, ignore_nan=False, **kw):
    return _j.dumps(obj, *a, **kw)


def dump(obj, fp, *a, ignore_nan=False, **kw):
    return _j.dump(obj, fp, *a, **kw)
