-- train.lua of deep-go, re-hosted: eval, eval_validation, train with the reference's bookkeeping
-- (EMA cost .95/.05, report every 10 iterations, validation + save every validation_interval, samples/s lines).
-- One eval per step: the reference's extra pcall'd eval (train.lua:106) only guarded against exceptions.
require 'data'
require 'dataloader'
local dg = require 'deepgo'
local ffi, lib = dg.ffi, dg.lib

function eval(model, grads, inputs, targets, criterion)
    grads:zero()
    local preds = model:forward(inputs, targets)
    local cost = criterion:forward(preds, targets)
    local df_do = criterion:backward(preds, targets)
    model:backward(inputs, df_do)
    return cost, grads
end

function eval_validation(experiment, input, targets)
    local model, batchSize = experiment.model, experiment.batchSize
    local validation_size = input.size
    local sum_costs, sum_errors = 0, 0
    for i = 1, validation_size / batchSize do
        local lo, hi = (i - 1) * batchSize + 1, math.min(i * batchSize, validation_size)
        local errors, cost = model:eval_slice(input, targets, lo, hi)     -- forward + top-1 + mean NLL on the device
        local size = hi - lo                                              -- range[2]-range[1], as train.lua:39
        sum_costs = sum_costs + cost * size
        sum_errors = sum_errors + errors
    end
    return sum_costs / validation_size, (1 - sum_errors / validation_size)
end

function train(experiment, params)
    local iters = params.iters
    local model, criterion, datasets = experiment.model, experiment.criterion, experiment.datasets
    local optimizer, batchSize = experiment.optimizer, experiment.batchSize
    local parameters, grads = experiment.modelParameters, experiment.grads
    local validationInput, validationOutput
    queue_on_minibatch(function(mb) validationInput, validationOutput = mb, mb.output end, datasets.validation, experiment.validationSize)
    do_queued_tasks()
    local validation_cost, validation_accuracy = -1, nil
    local train_costs, cost_average = {}, nil
    local total_start_time = os.clock()

    local function train_iter(train_set)
        local iter_start_time = os.clock()
        -- eval() and the update are ONE device call unless the reference does something between them (validation + save,
        -- train.lua:119-124): those iterations keep the separate calls, so every printed number is the reference's
        local fused = model.train_step ~= nil and (experiment.iterations + 1) % experiment.validation_interval ~= 0
        local ok, train_cost = pcall(function()
            if fused then return model:train_step(train_set, optimizer.rate) end
            local c = eval(model, grads, train_set, train_set.output, criterion)
            return c
        end)
        if not ok then global_input, global_output = train_set, train_set.output; error(train_cost) end
        local iter_time = os.clock() - iter_start_time
        cost_average = cost_average or train_cost
        cost_average = .95 * cost_average + .05 * train_cost
        experiment.iterations = experiment.iterations + 1
        if experiment.iterations % 10 == 0 then
            if experiment.iterations % experiment.validation_interval == 0 then
                validation_cost, validation_accuracy = eval_validation(experiment, validationInput, validationOutput)
                print("validation at iteration " .. experiment.iterations .. ": cost=" .. validation_cost .. ", accuracy=" .. validation_accuracy)
                table.insert(experiment.validation_costs, validation_cost)
                experiment:save()
            else
                print("training", cost_average, "(samples per second " .. batchSize / iter_time .. ")")
            end
            table.insert(train_costs, cost_average)
        end
        if fused then optimizer.rate = optimizer.rate * (1 - optimizer.rate_decay)      -- the decay of SGD:step (optimizer.lua:26)
        else optimizer:step(parameters, grads) end
    end

    for i = 1, iters do queue_on_minibatch(train_iter, datasets.train, batchSize) end
    do_queued_tasks()
    print("total samples per second", batchSize * iters / (os.clock() - total_start_time))
    return train_costs, experiment.validation_costs
end
