"""Small helpers shared by the host-side mirror (reference: torchao/utils.py)."""
import torch

__all__ = ["find_multiple", "fill_defaults", "get_model_size_in_bytes"]


def find_multiple(n: int, *args: int) -> int:
    """reference: utils.py:177-181"""
    import math
    from functools import reduce
    k = reduce(lambda x, y: x * y // math.gcd(x, y), args + (1,))
    return n if n % k == 0 else n + k - (n % k)


def fill_defaults(args, n, defaults_tail):
    """reference: utils.py:891-913"""
    if n - len(defaults_tail) > len(args):
        raise RuntimeError("not enough defaults to fill arguments")
    r = list(args)
    for i in range(len(args), n):
        r.append(defaults_tail[i - n + len(defaults_tail)])
    return r


def _tensor_bytes(t):
    if hasattr(t, "__tensor_flatten__"):
        names, _ = t.__tensor_flatten__()
        return sum(_tensor_bytes(getattr(t, n)) for n in names)
    return t.numel() * t.element_size()


def get_model_size_in_bytes(model, ignore_embeddings=False):
    """Bytes of parameters + buffers, recursing into quantized tensor subclasses.
    reference: utils.py:247-274 (the 'model size' of generate.py's bandwidth metric)."""
    size = 0
    for child in model.children():
        if not (isinstance(child, torch.nn.Embedding) and ignore_embeddings):
            for p in list(child.parameters(recurse=False)) + list(child.buffers(recurse=False)):
                size += _tensor_bytes(p)
            size += get_model_size_in_bytes(child, ignore_embeddings)
    return size
