"""Model drivers that exercise the collective end to end (ResNet-50 and GPT-2-medium DDP):
``zoo.build`` constructs the models, ``ddp_bench`` is the torchrun entry point."""
from .zoo import build  # noqa: F401
