"""Host mirror of optimizer.lua:16-27.  SGD.new(rate, decay); SGD:step(parameters, grad) applies
p += -rate*g on the device (dg_sgd_step) and keeps the multiplicative rate decay on the host,
applied after use exactly as optimizer.lua:26.  AdagradOptimizer is dead code in the reference
(reads an undefined global, never instantiated) and is not provided."""


class SGD:
    def __init__(self, rate, decay):
        self.rate = rate
        self.rate_decay = decay

    @staticmethod
    def new(rate, decay):
        return SGD(rate, decay)

    def step(self, parameters, grad):
        """`parameters` / `grad` are the model's flat-vector handles (model.getParameters())."""
        assert parameters.model is grad.model
        parameters.model.engine.sgd_step(self.rate)
        self.rate = self.rate * (1 - self.rate_decay)

    def after_fused_step(self):
        """The update itself ran inside dg_train_step with self.rate; what is left of SGD:step is the decay (optimizer.lua:26)."""
        self.rate = self.rate * (1 - self.rate_decay)
