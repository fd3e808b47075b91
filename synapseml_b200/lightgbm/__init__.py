from .native import lightgbmlib, LightGBMNativeError  # noqa: F401
