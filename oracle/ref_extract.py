"""Load the reference's own ``ArcMarginProduct`` at run time -- TEST INFRASTRUCTURE ONLY.

``/root/reference/ARCFACE/arc_main.py`` cannot be imported (argparse, NCCL init and dataset
globbing run at module scope, arc_main.py:34-48,:209-218; ``timm`` :24 is not installed), so the
``ClassDef`` at arc_main.py:130-176 is pulled out with ``ast`` and executed unchanged.  Nothing is
copied into this repository: the source is read where it lies.  ``forward`` hard-codes
``device='cuda'`` (arc_main.py:168); on a GPU-less machine :class:`CudaToInputDevice` rewrites that
keyword to the CPU so the body still runs verbatim.

``/root/reference`` exists only in the build container -- never on the GPU box -- so this module is
used solely by ``tests/golden/make_golden.py`` and by CPU tests that skip when the tree is absent.
"""
from __future__ import annotations

import ast
import os

REFERENCE_FILE = "/root/reference/ARCFACE/arc_main.py"


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_FILE)


def load_reference_class():
    """Return the reference ``ArcMarginProduct`` class object (arc_main.py:130-176)."""
    import torch  # noqa: F401  (needed by the exec'd namespace)

    with open(REFERENCE_FILE, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read())
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ArcMarginProduct")
    ns: dict = {}
    exec(  # noqa: S102 - the names the class body expects (arc_main.py:2-12,:19)
        "import math, torch\nimport torch.nn as nn\nimport torch.nn.functional as F\n"
        "from torch.nn import Parameter",
        ns,
    )
    exec(compile(ast.Module(body=[node], type_ignores=[]), REFERENCE_FILE, "exec"), ns)  # noqa: S102
    return ns["ArcMarginProduct"]


def cuda_to_input_device():
    """Context manager that maps the hard-coded ``device='cuda'`` (arc_main.py:168) to the CPU."""
    from torch.overrides import TorchFunctionMode

    class CudaToInputDevice(TorchFunctionMode):
        def __torch_function__(self, func, types, args=(), kwargs=None):
            kwargs = kwargs or {}
            if kwargs.get("device") == "cuda":
                kwargs = dict(kwargs, device="cpu")
            return func(*args, **kwargs)

    return CudaToInputDevice()


def reference_forward_backward(x, w, label, s=30.0, m=0.50, easy_margin=False):
    """Run reference head + ``nn.CrossEntropyLoss`` + ``loss.backward()`` on CPU tensors (fp32).

    Mirrors the call sequence at arc_main.py:326-329.  Returns dict(logits, loss, dx, dw).
    """
    import torch

    Ref = load_reference_class()
    x = x.detach().clone().float().requires_grad_(True)
    with cuda_to_input_device():
        head = Ref(w.shape[1], w.shape[0], s=s, m=m, easy_margin=easy_margin)
        with torch.no_grad():
            head.weight.copy_(w)
        out = head(x, label)
        loss = torch.nn.CrossEntropyLoss()(out, label.view(-1).long())
        loss.backward()
    return {
        "logits": out.detach(),
        "loss": float(loss.detach()),
        "dx": x.grad.detach(),
        "dw": head.weight.grad.detach(),
    }
