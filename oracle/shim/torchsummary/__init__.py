def summary(*a, **k):
    raise NotImplementedError
