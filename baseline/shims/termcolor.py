def colored(text, *a, **kw):
    return text
