"""InductionGraph: Encoder -> Induction -> Relation modules and the MSE loss - host mirror of the
reference's ``model/graph.py`` (``InductionGraph(N, K, Q, **kwds)``, ``forward()``, ``build_loss()``).

``forward()`` runs the fused device pipeline on the batch last fed (see ``model/base.py``):
embedding lookups (model/base.py:76-86) -> BiLSTM over all L steps with output dropout
(graph.py:34-49) -> structured self-attention pooling (graph.py:52-53, layer_module.py:49-98) ->
support/query split (graph.py:57-62) -> dynamic routing (graph.py:71-77, layer_module.py:101-150) ->
neural tensor layer (graph.py:85, layer_module.py:8-47), with loss and accuracy fused into the last
kernel.  ``build_loss`` exposes the mean-squared-error scalar; as in the reference the L2 term is
NOT part of the loss (graph.py:108 is overwritten by :109).
"""
from __future__ import annotations

from .base import Base


class InductionGraph(Base):
    def __init__(self, N, K, Q, **kwds):
        """
        N: Num of classes for each batch
        K: Num of instances for each class in the support set
        Q: Num of instances for each class in the query set
        """
        Base.__init__(self, kwds)
        self.num_classes = N
        self.support_num_per_class = K
        self.query_num_per_class = Q
        self.build()

    def forward(self):
        c = self.cfg
        train = getattr(self, "_train_mode", False)
        self.pipeline.forward(train=train)
        self.probs = self.pipeline.probs if c.B > 1 else self.pipeline.probs[0]
        self.alphas = self.pipeline.alphas.view(c.X, c.L, 1)
        self.class_vector = self.pipeline.cvec.view(c.B, c.N, c.H)
        self.support_encoder = self.pipeline.es
        self.query_encoder = self.pipeline.eq

    def build_loss(self):
        self.loss = float(self.pipeline.out2[0].item())
