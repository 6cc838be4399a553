def dispatch_command(*a, **k):
    raise RuntimeError("argh stub")
