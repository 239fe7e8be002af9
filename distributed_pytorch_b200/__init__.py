"""distributed_pytorch_b200: B200-native (sm_100a) data-parallel training of the VGG family on CIFAR-10-shaped data.

Same capabilities as BrianZCS/distributed_pytorch (single process, gather->scatter, per-parameter all-reduce, bucketed
DDP; same CLI, same model tree / state_dict), re-designed around hand-written tcgen05/TMA kernels, NVLink peer-memory
collectives and whole-step CUDA graphs.  See DESIGN.md.
"""
from .models.vgg import VGG, VGG11, VGG13, VGG16, VGG19
from .optim.sgd import SGD
from .trainer import CrossEntropyLoss, TrainStep, test_model, train_model
from .parallel.strategies import AllReduceSync, GatherScatterSync, NoSync
from .parallel.ddp import DistributedDataParallel

__version__ = "0.1.0"
__all__ = ["VGG", "VGG11", "VGG13", "VGG16", "VGG19", "SGD", "CrossEntropyLoss", "TrainStep", "train_model",
           "test_model", "AllReduceSync", "GatherScatterSync", "NoSync", "DistributedDataParallel"]
