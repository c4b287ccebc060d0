from .lenet import LeNet, LeNetConfig, MNIST_LENET, CIFAR_WIDE_LENET, loss_fn  # noqa: F401
