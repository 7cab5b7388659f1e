"""CUDA-graph replay of an inference forward (launch-bound shapes: DnCNN on one 256x256 image is 17 kernels of ~16 us).

The engines' launch lists are static per (batch, H, W): every pointer is plan workspace or a packed-operand buffer, so
one forward can be captured once and replayed with a single `cudaGraphLaunch`.  The programmatic-dependent-launch edges
between the persistent kernels are kept inside the graph.  Weight changes are honoured: the packed bf16 operands are
refreshed (eagerly, outside the graph) before a replay whenever the engine's change key moved."""
import torch


class GraphedInference:
    """y = GraphedInference(model, example)(x) == model(x) under torch.no_grad() for inputs of example's shape.

    The returned tensor is the graph's static output buffer: consume (or clone) it before the next call."""

    def __init__(self, model, example):
        if not example.is_cuda:
            raise RuntimeError("kair_b200 GraphedInference needs a CUDA example input (there is no CPU fallback)")
        self.model = model
        self.x = example.detach().clone()
        with torch.no_grad():
            side = torch.cuda.Stream(device=example.device)
            side.wait_stream(torch.cuda.current_stream(example.device))
            with torch.cuda.stream(side):          # warm-up: builds the engine / plan and packs the weights
                for _ in range(2):
                    model(self.x)
            torch.cuda.current_stream(example.device).wait_stream(side)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.y = model(self.x)
        self._key = self._change_key()

    def _engine(self):
        return self.model.engine(self.x.device)

    def _change_key(self):
        eng = self._engine()
        return eng._param_versions() if hasattr(eng, "_param_versions") else eng._versions()

    def __call__(self, x):
        if x.shape != self.x.shape:
            raise RuntimeError("kair_b200 GraphedInference was captured for shape %s, got %s"
                               % (tuple(self.x.shape), tuple(x.shape)))
        key = self._change_key()
        if key != self._key:
            with torch.no_grad():
                self.model(self.x)     # eager forward: re-packs the changed weights into the buffers the graph reads
            self._key = self._change_key()
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.y
