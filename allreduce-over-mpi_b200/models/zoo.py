"""Model constructors used by the DDP drivers (random init, no downloads).

BASELINE.json configs #4 and #5: ResNet-50 (torchvision) on synthetic ImageNet-shape batches and GPT-2
medium (transformers, 24 layers x 1024 x 16 heads, 355M parameters) with bf16 autocast. A tiny MLP and
a tiny GPT-2 exist so that the driver logic can be tested in seconds."""
import torch


def build(name, device, dtype=torch.float32):
    """Returns (model, make_batch(batch_size, generator) -> (inputs, targets), loss_fn, items_per_sample)."""
    if name == "resnet50":
        import torchvision
        model = torchvision.models.resnet50(weights=None).to(device).to(memory_format=torch.channels_last)

        def make_batch(bs, g):
            x = torch.randn(bs, 3, 224, 224, device=device, generator=g).contiguous(memory_format=torch.channels_last)
            y = torch.randint(0, 1000, (bs,), device=device, generator=g)
            return x, y

        def loss_fn(m, x, y):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = m(x)
            return torch.nn.functional.cross_entropy(out.float(), y)
        return model, make_batch, loss_fn, 1
    if name in ("gpt2-medium", "gpt2-tiny"):
        from transformers import GPT2Config, GPT2LMHeadModel
        if name == "gpt2-medium":
            cfg = GPT2Config(n_layer=24, n_head=16, n_embd=1024, n_positions=1024)
            seq = 1024
        else:
            cfg = GPT2Config(n_layer=2, n_head=4, n_embd=128, n_positions=128, vocab_size=1024)
            seq = 128
        cfg.attn_pdrop = cfg.embd_pdrop = cfg.resid_pdrop = 0.0
        model = GPT2LMHeadModel(cfg).to(device)
        if dtype != torch.float32:
            model = model.to(dtype)  # bf16 parameters -> bf16 gradients on the wire

        def make_batch(bs, g):
            ids = torch.randint(0, cfg.vocab_size, (bs, seq), device=device, generator=g)
            return ids, ids

        def loss_fn(m, x, y):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return m(input_ids=x, labels=y).loss
        return model, make_batch, loss_fn, seq
    if name == "mlp":
        model = torch.nn.Sequential(torch.nn.Linear(512, 2048), torch.nn.GELU(), torch.nn.Linear(2048, 2048),
                                    torch.nn.GELU(), torch.nn.Linear(2048, 100)).to(device)

        def make_batch(bs, g):
            return (torch.randn(bs, 512, device=device, generator=g),
                    torch.randint(0, 100, (bs,), device=device, generator=g))

        def loss_fn(m, x, y):
            return torch.nn.functional.cross_entropy(m(x), y)
        return model, make_batch, loss_fn, 1
    raise ValueError(f"unknown model {name}")
