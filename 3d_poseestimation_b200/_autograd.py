"""autograd bridge of the training path (filled in with the training kernels)."""


def lifter_apply(model, x, needs_grad):
    raise NotImplementedError("training path not built yet")
