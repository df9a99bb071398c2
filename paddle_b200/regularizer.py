"""paddle.regularizer (reference: python/paddle/regularizer.py)."""


class WeightDecayRegularizer:
    def __init__(self, coeff=0.0):
        self._coeff = float(coeff)
        self.coeff = self._coeff


class L2Decay(WeightDecayRegularizer):
    pass


class L1Decay(WeightDecayRegularizer):
    pass
