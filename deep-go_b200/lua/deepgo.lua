-- LuaJIT FFI binding of libdeepgo_b200.so.  The declarations are the text of include/deepgo.h itself
-- (preprocessor lines and the extern "C" braces removed), so this file cannot drift from the C ABI.
-- NOT EXECUTED in the build container (no LuaJIT there); the identical header text is exercised through
-- Python cffi by tests/test_abi.py, whose cdef parser has the same no-preprocessor constraint.
local ffi = require 'ffi'

local M = {}

local function script_dir()
    local src = debug.getinfo(1, 'S').source:sub(2)
    return src:match('(.*/)') or './'
end

M.root = os.getenv('DEEPGO_ROOT') or (script_dir() .. '../../')

local function header_cdef()
    local out = {}
    for line in io.lines(M.root .. 'include/deepgo.h') do
        local s = line:match('^%s*(.-)%s*$')
        if not (s:sub(1, 1) == '#' or s:sub(1, 10) == 'extern "C"' or s == '}') then
            out[#out + 1] = line
        end
    end
    return table.concat(out, '\n')
end

ffi.cdef(header_cdef())
M.lib = ffi.load(os.getenv('DEEPGO_LIB') or (M.root .. 'deep-go_b200/libdeepgo_b200.so'))
M.ffi = ffi
M.PREC_FP32, M.PREC_BF16 = 0, 1
M.FMT_F32_NCHW, M.FMT_BITPLANES = 0, 1

-- every entry point returns int; non-zero becomes a Lua error, so pcall() guards (train.lua:106) keep working
function M.check(rc)
    if rc ~= 0 then error(ffi.string(M.lib.dg_last_error()), 2) end
end

function M.context(device)
    local p = ffi.new('dg_ctx *[1]')
    M.check(M.lib.dg_ctx_create(device or 0, p))
    return ffi.gc(p[0], M.lib.dg_ctx_destroy)
end

-- numGPUs > 1 inside this process (experiments.lua:155-168): one context standing for devices 0 .. n-1
function M.context_multi(n)
    local devs = ffi.new('int[?]', n)
    for i = 0, n - 1 do devs[i] = i end
    local p = ffi.new('dg_ctx *[1]')
    M.check(M.lib.dg_ctx_create_multi(n, devs, p))
    return ffi.gc(p[0], M.lib.dg_ctx_destroy)
end

function M.model(ctx, numLayers, channelSize, kFirst, precision, maxBatch)
    local p = ffi.new('dg_model *[1]')
    M.check(M.lib.dg_model_create(ctx, numLayers, channelSize, 37, kFirst or 5, precision or M.PREC_BF16, maxBatch, p))
    return ffi.gc(p[0], M.lib.dg_model_destroy)
end

return M
