"""`import model as mdl; mdl.VGG11()` keeps working (reference main.py:12,101): re-export of the B200 model."""
from distributed_pytorch_b200.models.vgg import VGG, VGG11, VGG13, VGG16, VGG19  # noqa: F401
