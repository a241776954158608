"""Host mirror of train.lua: eval (:4-12), eval_validation (:14-45) and train (:47-142) with the reference's
bookkeeping (EMA cost .95/.05, print every 10 iterations, validation + save every validation_interval).
The reference's duplicate eval inside pcall (train.lua:106) is reduced to one eval."""
import time

import numpy as np

from . import data as _data


def eval(model, grads, inputs, targets, criterion):
    grads.zero()
    preds = model.forward(inputs, targets)
    cost = criterion.forward(preds, targets)
    df_do = criterion.backward(preds, targets)
    model.backward(inputs, df_do)
    return cost, grads


def eval_validation(experiment, input, targets):
    """train.lua:14-45, aggregate quirks included: `for i=1,V/batchSize` drops the tail, size = batchSize-1
    (range[2]-range[1]), divisor = V."""
    model, criterion, batchSize = experiment.model, experiment.criterion, experiment.batchSize
    validation_size = input.shape[0]
    sum_costs, sum_errors = 0.0, 0
    for i in range(1, int(validation_size // batchSize) + 1):
        lo, hi = (i - 1) * batchSize, min(i * batchSize, validation_size)
        batch, batch_targets = input[lo:hi], targets[lo:hi]
        preds = model.forward(batch, batch_targets)
        _, top_preds = preds.max(2)
        errors = int((top_preds != batch_targets.astype(np.int64)).sum())
        cost = criterion.forward(preds, batch_targets)
        size = hi - (lo + 1)
        sum_costs += cost * size
        sum_errors += errors
    return sum_costs / validation_size, 1 - sum_errors / validation_size


def train(experiment, params):
    iters = params["iters"]
    model, criterion, datasets = experiment.model, experiment.criterion, experiment.datasets
    optimizer, batchSize = experiment.optimizer, experiment.batchSize
    parameters, grads = experiment.modelParameters, experiment.grads
    base = params.get("base", "")
    state = {"validation": None, "cost_average": None, "validation_cost": -1}
    train_costs = []

    def set_validation_data(minibatch):
        state["validation"] = (minibatch["input"], minibatch["output"])

    _data.queue_on_minibatch(set_validation_data, datasets["validation"], experiment.validationSize, base)
    _data.do_queued_tasks()
    total_start = time.time()

    # eval() + optimizer:step are ONE device call (dg_train_step: captured step graph, in-graph gradient exchange on several GPUs)
    # whenever nothing has to happen between them.  The reference validates and saves BETWEEN eval and the update
    # (train.lua:111-131): those iterations keep the two separate calls, so every printed number is what the reference prints.
    fused_ok = hasattr(model, "train_step") and hasattr(optimizer, "after_fused_step") and hasattr(criterion, "fused") \
        and not params.get("separate_calls", False)

    def train_iter(train_set):
        t0 = time.time()
        fused = fused_ok and (experiment.iterations + 1) % experiment.validation_interval != 0
        if fused:
            train_cost = model.train_step(train_set["input"], train_set["output"], optimizer.rate)
        else:
            train_cost, _ = eval(model, grads, train_set["input"], train_set["output"], criterion)
        iter_time = time.time() - t0
        if state["cost_average"] is None:
            state["cost_average"] = train_cost
        state["cost_average"] = .95 * state["cost_average"] + .05 * train_cost
        experiment.iterations += 1
        if experiment.iterations % 10 == 0:
            if experiment.iterations % experiment.validation_interval == 0:
                vc, va = eval_validation(experiment, *state["validation"])
                state["validation_cost"] = vc
                print("validation at iteration %d: cost=%s, accuracy=%s" % (experiment.iterations, vc, va))
                experiment.validation_costs.append(vc)
                experiment.save()
            else:
                print("training", state["cost_average"], "(samples per second %s)" % (batchSize / iter_time))
            train_costs.append(state["cost_average"])
        if fused:
            optimizer.after_fused_step()
        else:
            optimizer.step(parameters, grads)

    for _ in range(iters):
        _data.queue_on_minibatch(train_iter, datasets["train"], batchSize, base)
    _data.do_queued_tasks()
    total_time = time.time() - total_start
    print("total samples per second", batchSize * iters / total_time)
    return train_costs, experiment.validation_costs
