"""Empty matplotlib stub (reference imports it at module scope; plotting is out of scope)."""
def use(*a, **k):
    pass
