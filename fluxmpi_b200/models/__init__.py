"""Model definitions for the benchmark configs named in BASELINE.json (the reference itself
ships none — model code lives in Lux/Flux user scripts)."""
from .mlp import quickstart_mlp, functional_forward
from .resnet import resnet50, resnet_tiny, ResNet
from .vit import vit_b16, vit_tiny, ViT
from .gpt2 import gpt2_medium, gpt2_tiny, GPT2
