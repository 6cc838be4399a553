"""Pass-through stand-in for gin-config, used ONLY by oracle/make_golden.py to
import the reference (/root/reference) in a container that has no gin.
Test infrastructure; never imported by the product package."""
REQUIRED = object()


def configurable(*args, **kwargs):
    if len(args) == 1 and callable(args[0]) and not kwargs:
        return args[0]

    def deco(fn):
        return fn
    return deco


def external_configurable(fn, *a, **k):
    return fn


def parse_config_file(*a, **k):
    return None


def operative_config_str():
    return ""
