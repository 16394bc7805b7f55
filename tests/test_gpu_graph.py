"""The head's whole step captured into one CUDA graph (graphed.GraphedHeadStep) replays to the eager results."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(256, 128, 5000), (640, 512, 70000)])  # second shape: normalisation inside the GEMM
def test_graphed_step_equals_eager(shape):
    from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
    from ddp_classification_pytorch_b200.graphed import GraphedHeadStep
    B, D, C = shape
    torch.manual_seed(3)
    head = ArcMarginProduct(D, C).cuda()
    batches = [(torch.randn(B, D, device="cuda"), torch.randint(0, C, (B,), device="cuda")) for _ in range(3)]
    eager = []
    for x, y in batches:
        head.weight.grad = None
        xg = x.clone().requires_grad_(True)
        loss = head.forward_loss(xg, y)
        loss.backward()
        eager.append((loss.item(), xg.grad.clone(), head.weight.grad.clone()))
        del loss, xg  # no autograd graph of the eager steps may outlive them (see GraphedHeadStep)
    head.weight.grad = None
    step = GraphedHeadStep(head, B)
    for (x, y), (l_e, dx_e, dw_e) in zip(batches, eager):
        loss = step(x, y)
        torch.cuda.synchronize()
        assert loss.item() == pytest.approx(l_e, rel=1e-6)
        torch.testing.assert_close(step.x.grad, dx_e, rtol=1e-3, atol=1e-8)      # fp32 atomics: summation order
        torch.testing.assert_close(head.weight.grad, dw_e, rtol=1e-3, atol=1e-8)
    _lib.check(_lib.lib().arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))


def test_graphed_step_rejects_cpu_head():
    from ddp_classification_pytorch_b200 import ArcMarginProduct
    from ddp_classification_pytorch_b200.graphed import GraphedHeadStep
    with pytest.raises(RuntimeError, match="CUDA"):
        GraphedHeadStep(ArcMarginProduct(16, 10), 4)
