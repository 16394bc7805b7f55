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


def test_graphed_step_with_cached_head_follows_weight_updates():
    """A head with the normalised-weight cache (cache_normalized_weight=True) stepped by torch.optim.SGD between
    replays: the captured step must re-normalise the CURRENT weight on every replay, not freeze the cache's
    'still valid' decision of capture time."""
    from ddp_classification_pytorch_b200 import ArcMarginProduct
    from ddp_classification_pytorch_b200.graphed import GraphedHeadStep
    B, D, C = 128, 64, 2000
    torch.manual_seed(5)
    w0 = (torch.rand(C, D) * 2 - 1) * 0.05
    batches = [(torch.randn(B, D, device="cuda"), torch.randint(0, C, (B,), device="cuda")) for _ in range(4)]

    def run(graphed):
        head = ArcMarginProduct(D, C, cache_normalized_weight=graphed).cuda()
        with torch.no_grad():
            head.weight.copy_(w0)
        opt = torch.optim.SGD(head.parameters(), lr=5.0)
        losses = []
        if graphed:
            with torch.no_grad():
                head.forward_loss(*batches[0])  # fills the cache: an eager forward would now skip the normalisation
            step = GraphedHeadStep(head, B)
        for x, y in batches:
            if graphed:
                loss = step(x, y)
            else:
                head.weight.grad = None
                loss = head.forward_loss(x, y)
                loss.backward()
            losses.append(loss.item())
            opt.step()
        torch.cuda.synchronize()
        return losses, head.weight.detach().clone()

    eager_losses, eager_w = run(False)
    graph_losses, graph_w = run(True)
    assert eager_losses[0] != pytest.approx(eager_losses[-1], rel=1e-3)  # the weight really moved
    assert graph_losses == pytest.approx(eager_losses, rel=1e-5)
    torch.testing.assert_close(graph_w, eager_w, rtol=1e-4, atol=1e-7)


def test_graphed_step_rejects_cpu_head():
    from ddp_classification_pytorch_b200 import ArcMarginProduct
    from ddp_classification_pytorch_b200.graphed import GraphedHeadStep
    with pytest.raises(RuntimeError, match="CUDA"):
        GraphedHeadStep(ArcMarginProduct(16, 10), 4)
