from .model import Model  # noqa: F401
from .net import Net  # noqa: F401
from .loss import Loss, SoftmaxCrossEntropyLoss, MSELoss  # noqa: F401
from .learningrate import LearningRate  # noqa: F401
from .input import (Dataloader, TFRecordDataLoader, DataPathDataLoader, PlaceholderDataLoader,  # noqa: F401
                    InputOptions, SyntheticDataLoader)
