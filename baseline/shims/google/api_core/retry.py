class Retry:
    def __init__(self, *a, **kw):
        pass

    def __call__(self, fn):
        return fn


def if_exception_type(*a, **kw):
    return lambda exc: False
