from .cifar10_cnn import Cifar10CNN, Cifar10SyntheticLoader  # noqa: F401
from .resnet import ResNet18, ResNet50, ImageNetSyntheticLoader  # noqa: F401
from .dehaze_fcn import DehazeFCN, DehazeSyntheticLoader  # noqa: F401
