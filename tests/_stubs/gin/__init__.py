"""Test-only stand-in for `gin-config` (absent from this image): just enough for the
reference modules to import. Not product code."""
REQUIRED = object()


def configurable(*args, **kwargs):
    if len(args) == 1 and callable(args[0]) and not kwargs:
        return args[0]

    def deco(fn):
        return fn

    return deco


def external_configurable(obj, *args, **kwargs):
    return obj


def parse_config_file(*args, **kwargs):
    raise RuntimeError("gin stub: no config parsing in tests")
