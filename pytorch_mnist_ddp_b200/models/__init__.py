from .net import Net, PARAM_SPECS, ParamLayout  # noqa: F401
