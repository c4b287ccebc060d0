from .flat import FlatOptimizer, OptHyper, StepLR, OPT_ADADELTA, OPT_SGD, OPT_ADAM  # noqa: F401
