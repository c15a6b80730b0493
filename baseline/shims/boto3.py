def client(*a, **kw):
    raise RuntimeError("boto3 is not installed")


def resource(*a, **kw):
    raise RuntimeError("boto3 is not installed")
