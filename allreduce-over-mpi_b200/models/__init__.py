"""Model drivers that exercise the collective end to end (ResNet-50 and GPT-2-medium DDP)."""
