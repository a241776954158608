-- localtest.lua: the reference's smoke configuration (3 layers x 64 filters, batch 2, 20 iterations on data/).
-- useCuda=false in the reference meant "CPU Torch7"; this port has no CPU path, the flag only skips makeDataParallel.
require 'experiments'

experiment = basicGoExperiment:new{
    useCuda = false,
    data_root = "data",
    validation_interval = 20,
    validationSize = 100,
    batchSize = 2
}
experiment:run{iters = 20}
