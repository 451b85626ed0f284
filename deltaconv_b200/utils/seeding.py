"""Name-keyed deterministic parameter fill, so that the reference model (in the build
container), the oracle and the CUDA product get bit-identical weights without shipping a
checkpoint: the value of every state-dict entry depends only on its key, shape and `seed`."""
import zlib

import torch


def fill_deterministic(module, seed=1):
    state = module.state_dict()
    with torch.no_grad():
        for key in sorted(state):
            t = state[key]
            if not t.is_floating_point():
                continue
            gen = torch.Generator().manual_seed((zlib.crc32(key.encode()) ^ (seed * 2654435761)) & 0x7FFFFFFF)
            leaf = key.rsplit('.', 1)[-1]
            if t.dim() >= 2:                                  # Linear weight
                val = torch.randn(t.shape, generator=gen) / (t.size(1) ** 0.5)
            elif leaf == 'weight':                            # BN gamma
                val = 1 + 0.1 * torch.randn(t.shape, generator=gen)
            elif leaf == 'running_var':
                val = 1 + 0.1 * torch.rand(t.shape, generator=gen)
            else:                                             # biases, BN beta, running_mean
                val = 0.1 * torch.randn(t.shape, generator=gen)
            t.copy_(val.to(t.dtype))
    return module
