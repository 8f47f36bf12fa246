# B200Ext.jl -- the package extension a DINCAE.jl maintainer adds to bind libdincae_b200.so (an `ext/` module next to
# ext/CUDAExt.jl, loaded with `using DINCAE, B200Ext`).
#
# NOT EXECUTED IN THIS REPOSITORY: the build image has no Julia runtime (and no network to install one), so this file is
# written by inspection, call site by call site, against the reference (gher-uliege/DINCAE.jl v2.0.6) and
# include/dincae_b200.h.  The same symbols, in the same order, are driven by the Python ctypes binding
# (dincae.jl_b200/lib.py, api.py) and by the plain-C harness tests/c_abi_smoke.c, which is what the test-suite runs.
#
# Usage -- the reference's drivers, unchanged:
#     DINCAE.reconstruct(B200Array{Float32}, data_all, fnames_rec; kw...)
#     DINCAE.reconstruct_points(Float32, B200Array{Float32}, filename, varname, grid, fname_rec; kw...)
# Every call the two drivers make on `Atype`, `device`, the model and the mini-batches is answered by a method below;
# INTEGRATION.md lists them line by line (src/model.jl:479-651, src/points.jl:541-655).
module B200Ext

import DINCAE
import DINCAE: interpnd!, interp_adjn!, _to_device, train_init, train_epoch!, loss_function, StepModel, PointCloud, sizex
import Flux
using Random

export B200Array

const lib = "libdincae_b200"
const HOST = Int32(0)
const DEVICE_XIN = Int32(2)

check(st::Integer) = st == 0 || error(unsafe_string(ccall((:dincae_last_error, lib), Cstring, ())))

# ---- the Atype ----------------------------------------------------------------------------------------------------
# `Atype(inputs_)` (src/data.jl:858, src/points.jl:338) must give an array the caller owns: DataBatches reuses its batch buffer
# and PrefetchDataIter fills it for the next mini-batch on another thread while this one trains.  B200Array is therefore a thin
# OWNING wrapper around a host Array (the copy the reference makes for Atype = Array as well); the library stages it to the GPU
# (dincae_batch_upload, one batch ahead).  It is an AbstractArray, so size / sum / Array(x) / extrema of the drivers' @info
# lines work unchanged.
struct B200Array{T,N} <: AbstractArray{T,N}
    data::Array{T,N}
end
B200Array{T}(a::AbstractArray{S,N}) where {T,S,N} = B200Array{T,N}(Array{T,N}(a))          # copies (also materialises views)
B200Array{T,N}(a::AbstractArray{S,N}) where {T,S,N} = B200Array{T,N}(Array{T,N}(a))
B200Array(a::AbstractArray{T,N}) where {T,N} = B200Array{T,N}(Array{T,N}(a))
Base.size(a::B200Array) = size(a.data)
Base.getindex(a::B200Array, i::Int...) = getindex(a.data, i...)
Base.setindex!(a::B200Array, v, i::Int...) = setindex!(a.data, v, i...)
Base.IndexStyle(::Type{<:B200Array}) = IndexLinear()
Base.similar(a::B200Array, ::Type{T}, dims::Dims) where {T} = B200Array(Array{T}(undef, dims))
Base.Array(a::B200Array) = a.data
Base.pointer(a::B200Array) = pointer(a.data)
Base.unsafe_convert(::Type{Ptr{T}}, a::B200Array{T}) where {T} = pointer(a.data)
host(a::B200Array) = a.data
host(a::Array) = a
host(a::AbstractArray) = Array(a)

# ---- handles ------------------------------------------------------------------------------------------------------
mutable struct Ctx
    h::Ptr{Cvoid}
    function Ctx(device::Integer = 0)
        r = Ref{Ptr{Cvoid}}(C_NULL)
        check(ccall((:dincae_ctx_create, lib), Int32, (Int32, Ref{Ptr{Cvoid}}), device, r))
        c = new(r[])
        finalizer(c -> ccall((:dincae_ctx_destroy, lib), Int32, (Ptr{Cvoid},), c.h), c)
    end
end
const default_ctx = Ref{Union{Nothing,Ctx}}(nothing)
ctx() = something(default_ctx[], (default_ctx[] = Ctx(parse(Int, get(ENV, "DINCAE_B200_DEVICE", "0"))); default_ctx[]))

# layout-compatible with `dincae_config` (include/dincae_b200.h): 35 4-byte fields, no padding
struct Config
    W::Int32; H::Int32; ninput::Int32; noutput::Int32; nlevels::Int32
    nfilter::NTuple{8,Int32}
    skip_mask::UInt32
    upsampling::Int32
    nchains::Int32
    loss_weights::NTuple{4,Float32}
    gamma::Float32
    truth_uncertain::Int32
    l1_beta::Float32; l2_beta::Float32
    laplacian_penalty::Float32; laplacian_error_penalty::Float32
    clip_grad::Float32
    adam_beta1::Float32; adam_beta2::Float32; adam_eps::Float32
    leaky_slope::Float32
    precision::Int32
    max_batch::Int32
end
pad8(v) = ntuple(i -> i <= length(v) ? Int32(v[i]) : Int32(0), 8)
pad4(v) = ntuple(i -> i <= length(v) ? Float32(v[i]) : 0f0, 4)

# ---- the model ----------------------------------------------------------------------------------------------------
# `model |> device` (src/model.jl:559, src/points.jl:577) receives the Flux StepModel that genmodel built on the CPU.  device()
# reads the topology off it, creates the library model lazily (max_batch is only known at the first mini-batch) and copies
# Flux.trainables in.  The Flux model is kept: Flux.state(cpu(model)) (src/model.jl:603) copies the trained weights back into it.
mutable struct B200Model
    flux::StepModel              # the CPU model genmodel returned (topology, hyper-parameters, checkpoint shell)
    h::Ptr{Cvoid}                # dincae_model*, C_NULL until the first call
    ctx::Ctx
    max_batch::Int
    clip_grad::Float32
    precision::Int32
end

"All Conv layers of one recmodel4 chain in Flux.trainables order (enc 1..L, dec L..1) -- src/model.jl:242-284"
function convs(chain)
    out = Flux.Conv[]
    walk(l::Flux.Conv) = push!(out, l)
    walk(l::Flux.Chain) = foreach(walk, l.layers)
    walk(l::Flux.SkipConnection) = walk(l.layers)
    walk(l) = nothing                                   # MeanPool, Upsample, croppadding closures, identity
    walk(chain)
    return out
end

"Levels whose recmodel4 block is wrapped in a SkipConnection (src/model.jl:277-279), as the 1-based list genmodel received"
function skiplevels(chain, L)
    lev = Int[]
    function walk(l, depth)
        if l isa Flux.SkipConnection
            push!(lev, depth); walk(l.layers, depth)
        elseif l isa Flux.Chain
            nconv = 0
            for x in l.layers
                x isa Flux.Conv && (nconv += 1)
                # the inner level sits between the encoder conv and the decoder conv of this level
                (x isa Flux.Chain || x isa Flux.SkipConnection) && walk(x, depth + 1)
            end
        end
    end
    walk(chain, 1)
    return lev
end

function config_of(m::StepModel, sz, max_batch, clip_grad, precision)
    cs = convs(m.chains[1])
    L = length(cs) ÷ 2
    ninput = size(cs[1].weight, 3)
    nfilter = [size(cs[l].weight, 4) for l in 1:L]
    nout2 = size(cs[end].weight, 4)
    mask = UInt32(0)
    for s in skiplevels(m.chains[1], L)
        1 <= s <= L && (mask |= UInt32(1) << (s - 1))
    end
    up = any(l -> l isa Flux.Upsample && l.mode == :bilinear, Flux.modules(m.chains[1])) ? 1 : 0
    # final_layer / costfun are closures over gamma, truth_uncertain and the penalties (src/model.jl:160-171): read the captures
    gamma = Float32(m.final_layer.gamma)
    cf = m.costfun
    length(m.chains) <= 2 || error("B200Ext: the reference builds at most two chains (src/model.jl:322-340)")
    return Config(sz[1], sz[2], ninput, nout2 ÷ 2, L, pad8(nfilter), mask, up, length(m.chains), pad4(m.loss_weights), gamma,
                  Int32(cf.truth_uncertain), Float32(m.regularization_L1), Float32(m.regularization_L2),
                  Float32(cf.laplacian_penalty), Float32(cf.laplacian_error_penalty), clip_grad,
                  0.9f0, 0.999f0, 1f-8, 0f0, precision, max_batch)
end

flatparams(m::StepModel) = reduce(vcat, [vec(Array{Float32}(p)) for p in Flux.trainables(m)])      # the order the library documents

function ensure!(m::B200Model, xin)
    B = size(xin)[end]
    if m.h == C_NULL || B > m.max_batch
        w = m.h == C_NULL ? flatparams(m.flux) : getparams(m)
        m.h != C_NULL && check(ccall((:dincae_model_destroy, lib), Int32, (Ptr{Cvoid},), m.h))
        m.max_batch = max(B, m.max_batch)
        cfg = config_of(m.flux, size(xin), m.max_batch, m.clip_grad, m.precision)
        r = Ref{Ptr{Cvoid}}(C_NULL)
        check(ccall((:dincae_model_create, lib), Int32, (Ptr{Cvoid}, Ref{Config}, Ref{Ptr{Cvoid}}), m.ctx.h, Ref(cfg), r))
        m.h = r[]
        check(ccall((:dincae_model_set_params, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int64), m.h, w, length(w)))
    end
    return m
end

function getparams(m::B200Model)
    n = Ref{Int64}(0)
    check(ccall((:dincae_model_num_params, lib), Int32, (Ptr{Cvoid}, Ref{Int64}), m.h, n))
    w = Vector{Float32}(undef, n[])
    check(ccall((:dincae_model_get_params, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int64), m.h, w, n[]))
    return w
end

"`device = _to_device(Atype)` (src/model.jl:485): identity on arrays, StepModel -> library model"
struct Device end
(::Device)(x::AbstractArray) = x
(::Device)(x::Nothing) = nothing
(::Device)(x::AbstractVector{<:NTuple}) = x                     # observation positions stay host vectors (src/points.jl:338)
function (::Device)(m::StepModel)
    bm = B200Model(m, C_NULL, ctx(), 0, 5f0, Int32(1))
    finalizer(b -> b.h != C_NULL && ccall((:dincae_model_destroy, lib), Int32, (Ptr{Cvoid},), b.h), bm)
    return bm
end
Base.show(io::IO, ::Device) = print(io, "B200 (libdincae_b200)")
@inline _to_device(::Type{<:B200Array}) = Device()

"`Flux.state(cpu(model))` at the save epochs (src/model.jl:603): the trained weights copied back into the Flux model"
function Flux.cpu(m::B200Model)
    m.h == C_NULL && return m.flux
    w = getparams(m); off = 0
    for p in Flux.trainables(m.flux)
        copyto!(p, reshape(view(w, off+1:off+length(p)), size(p))); off += length(p)
    end
    return m.flux
end

"model(xin): (W,H,ninput,B) -> (W,H,2*noutput,B) = (mean, sigma^2)  (StepModel call, src/model.jl:182-189, :561, :627)"
function (m::B200Model)(xin::AbstractArray{Float32,4})
    ensure!(m, xin)
    x = host(xin)
    out = Array{Float32}(undef, size(x, 1), size(x, 2), size(convs(m.flux.chains[1])[end].weight, 4), size(x, 4))
    check(ccall((:dincae_forward, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int32, Ptr{Float32}, Int32), m.h, x, size(x, 4), out, HOST))
    return B200Array(out)
end

"loss_function(model, xin, xtrue) on gridded truth (src/model.jl:192-222, :567)"
function loss_function(m::B200Model, xin::AbstractArray{Float32,4}, xtrue::AbstractArray{Float32,4})
    ensure!(m, xin)
    l = Ref{Float64}()
    check(ccall((:dincae_loss, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Int32, Int32, Ref{Float64}),
                m.h, host(xin), host(xtrue), size(xin, 4), HOST, l))
    return l[]
end

# point-cloud truth: Vector of (pos = Vector{NTuple{2,Float32}}, x = (P,2) array) per sample (src/points.jl:317-340)
const PointTruth = AbstractVector{<:NamedTuple{(:pos, :x)}}
function pack_truth(xtrue::PointTruth)
    offsets = Int64[0]
    for xt in xtrue
        push!(offsets, offsets[end] + length(xt.pos))
    end
    P = offsets[end]
    pos = Vector{NTuple{2,Float32}}(undef, P)            # interleaved (x,y) pairs in memory = the ABI's (P,2) layout
    x = Matrix{Float32}(undef, P, 2)                     # column-major: values, then 1/err var
    for (b, xt) in enumerate(xtrue)
        r = offsets[b]+1:offsets[b+1]
        pos[r] .= xt.pos
        x[r, :] .= host(xt.x)
    end
    return pos, x, offsets
end

"loss_function(model, xin, xtrue) on point-cloud truth (src/points.jl:354-426, :593)"
function loss_function(m::B200Model, xin::AbstractArray{Float32,4}, xtrue::PointTruth)
    ensure!(m, xin)
    pos, x, offsets = pack_truth(xtrue)
    l = Ref{Float64}()
    check(ccall((:dincae_loss_points, lib), Int32,
                (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ptr{Int64}, Int32, Int32, Ref{Float64}),
                m.h, host(xin), pos, x, offsets, size(xin, 4), HOST, l))
    return l[]
end

"train_init(model, :ADAM; clip_grad, learning_rate) (src/flux.jl:7-21, src/model.jl:585): Adam + ClipGrad state lives in the library"
function train_init(m::B200Model, optim; clip_grad = nothing, learning_rate = nothing)
    @assert optim == :ADAM
    m.clip_grad = isnothing(clip_grad) ? 0f0 : Float32(clip_grad)
    m.h != C_NULL && (w = getparams(m); check(ccall((:dincae_model_destroy, lib), Int32, (Ptr{Cvoid},), m.h)); m.h = C_NULL;
                      copyparams!(m.flux, w))           # rebuilt with the clip value at the first training batch
    return (m, nothing, nothing)
end
function copyparams!(flux, w)
    off = 0
    for p in Flux.trainables(flux)
        copyto!(p, reshape(view(w, off+1:off+length(p)), size(p))); off += length(p)
    end
end

"""
train_epoch!((model,params,opt_state), dl, learning_rate; clip_grad) (src/flux.jl:24-40; called at src/model.jl:592 with
PrefetchDataIter(train) and at src/points.jl:620 with the PointCloud): mean of the per-batch losses.  Gridded batches go
through the two upload slots, so the host->device copy of batch i+1 overlaps the kernels of batch i.
"""
function train_epoch!((m, _, _)::Tuple{B200Model,Nothing,Nothing}, dl, learning_rate; clip_grad = nothing)
    loss_sum = 0.0; N = 0; loss = Ref{Float32}()
    it = iterate(dl)
    it === nothing && error("empty data iterator")
    (xin, xtrue), state = it
    if xtrue isa PointTruth                                                # reconstruct_points
        while true
            ensure!(m, xin)
            pos, x, offsets = pack_truth(xtrue)
            check(ccall((:dincae_train_step_points, lib), Int32,
                        (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ptr{Int64}, Int32, Float32, Int32, Ref{Float32}),
                        m.h, host(xin), pos, x, offsets, size(xin, 4), learning_rate, HOST, loss))
            loss_sum += loss[]; N += 1
            it = iterate(dl, state)
            it === nothing && break
            (xin, xtrue), state = it
        end
        return loss_sum / N
    end
    ensure!(m, xin)
    slot = Int32(0)
    upload(s, a, b) = check(ccall((:dincae_batch_upload, lib), Int32, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Int32),
                                  m.h, s, host(a), host(b), size(a, 4)))
    upload(slot, xin, xtrue)
    while true
        nxt = iterate(dl, state)
        xin2 = xtrue2 = nothing
        if nxt !== nothing
            (xin2, xtrue2), state = nxt
            upload(slot ⊻ Int32(1), xin2, xtrue2)
        end
        GC.@preserve xin xtrue xin2 xtrue2 begin
            check(ccall((:dincae_train_step_slot, lib), Int32, (Ptr{Cvoid}, Int32, Int32, Float32, Ref{Float32}),
                        m.h, slot, size(xin, 4), learning_rate, loss))
        end
        loss_sum += loss[]; N += 1
        nxt === nothing && break
        xin, xtrue, slot = xin2, xtrue2, slot ⊻ Int32(1)
    end
    return loss_sum / N
end

# ---- the two kernels of ext/CUDAExt.jl, same 3-argument signatures, dispatched on the array type ------------------------
"interpnd!(pos, A, vec) (src/points.jl:6-27, ext/CUDAExt.jl:31): vec .+= bilinear gather of A at pos"
function interpnd!(pos::AbstractVector{NTuple{2,Float32}}, A::B200Array{Float32,2}, vec)
    tmp = Vector{Float32}(undef, length(pos))
    check(ccall((:dincae_interp, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int64, Ptr{Float32}, Int32, Int32, Ptr{Float32}, Int32),
                ctx().h, pos, length(pos), A.data, size(A, 1), size(A, 2), tmp, HOST))
    vec .+= tmp                          # the reference accumulates into vec (interpnd zeroes it first, src/points.jl:30-34)
    return nothing
end

"interp_adjn!(pos, values, B) (src/points.jl:52-71, ext/CUDAExt.jl:71): B .= adjoint scatter; sorted + segmented, no float atomics"
function interp_adjn!(pos::AbstractVector{NTuple{2,Float32}}, values::B200Array{Float32,1}, B)
    out = Matrix{Float32}(undef, size(B, 1), size(B, 2))
    check(ccall((:dincae_interp_adjoint, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Int64, Int32, Int32, Ptr{Float32}, Int32),
                ctx().h, pos, values.data, length(pos), size(B, 1), size(B, 2), out, HOST))
    B .= out
    return nothing
end

"""vector2_costfun(xrec, xtrue, truth_uncertain, directionobs) (src/vector2.jl:53-100) on B200Arrays: the cost of the 2-D vector
model (output_ndims = 2) and, for its pullback, the gradient w.r.t. xrec from the same call."""
function vector2_costfun_and_grad(xrec::B200Array{Float32,4}, xtrue::B200Array{Float32,4}, truth_uncertain::Bool, directionobs)
    W, H, five, B = size(xrec)
    nsites = size(directionobs, 3)
    (five == 5 && size(xtrue) == (W, H, 2nsites, B)) || error("vector2_costfun: sizes do not fit")
    cost = Ref{Float64}(0)
    grad = Array{Float32}(undef, W, H, 5, B)
    d = directionobs isa B200Array ? directionobs.data : Array{Float32}(directionobs)
    check(ccall((:dincae_vector2_loss, lib), Int32,
                (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int32, Int32, Int32, Int32, Int32, Int32, Int32, Ref{Float64}, Ptr{Float32}),
                ctx().h, xrec.data, xtrue.data, d, W, H, nsites, B, ndims(directionobs) == 4, truth_uncertain, HOST, cost, grad))
    return Float32(cost[]), grad
end
DINCAE.vector2_costfun(xrec::B200Array{Float32,4}, xtrue::B200Array{Float32,4}, truth_uncertain, directionobs) =
    vector2_costfun_and_grad(xrec, xtrue, truth_uncertain, directionobs)[1]

# ---- reconstruct_points: mini-batch gridding on the device ---------------------------------------------------------------
# Base.iterate(::PointCloud) (src/points.jl:317-340) grids 2*ntime_win scatters per sample with the CPU loop of interp_adjn!.
# For Atype <: B200Array the same mini-batch comes from dincae_points_assemble: the dataset (positions after posindices!
# filtering, values, aux cube) is uploaded once, the track dropout (integer work, :284-300) is drawn here, position jitter and
# the sorted scatters run on the GPU.
const datasets = IdDict{Any,Ptr{Cvoid}}()
function handle(d::PointCloud{<:B200Array,Float32,2})
    get!(datasets, d) do
        N = length(d.x)
        offsets = Int64[0]; for p in d.pos; push!(offsets, offsets[end] + length(p)); end
        pos = reduce(vcat, d.pos); x = reduce(vcat, d.x)
        naux = size(d.auxdata, 3)
        r = Ref{Ptr{Cvoid}}(C_NULL)
        check(ccall((:dincae_points_create, lib), Int32,
                    (Ptr{Cvoid}, Int32, Int32, Int64, Ptr{Float32}, Ptr{Float32}, Ptr{Int64}, Ptr{Float32}, Int32,
                     Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int32, Float32, Float32, Ref{Ptr{Cvoid}}),
                    ctx().h, length(d.grid_scaled[1]), length(d.grid_scaled[2]), N, pos, x, offsets,
                    naux > 0 ? d.auxdata : C_NULL, naux, d.grid_scaled[1], d.grid_scaled[2], d.dayofyear_cos, d.dayofyear_sin,
                    d.ntime_win, d.jitter_std_pos[1], d.jitter_std_pos[2], r))
        r[]
    end
end

function Base.iterate(d::PointCloud{Atype,Float32,2}, index = 0) where {Atype<:B200Array}
    ntime = length(d.x)
    bs = index+1:min(index + d.batch_size, ntime)
    isempty(bs) && return nothing
    sz = sizex(d)
    idx = Int64[d.perm[j] - 1 for j in bs]                                 # 0-based time indices
    keep = UInt8[]; koff = Int64[0]
    for j in bs                                                            # track dropout of the centre step (src/points.jl:284-300)
        ids = d.id[d.perm[j]]
        sel = trues(length(ids))
        if d.train
            uid = unique(ids)
            for u in uid
                rand() < d.probability_skip_for_training && (sel[ids .== u] .= false)
            end
            sum(sel) == 0 && !isempty(uid) && (sel[ids .== rand(uid)] .= true)
        end
        append!(keep, UInt8.(sel)); push!(koff, length(keep))
    end
    xin = Array{Float32}(undef, sz..., length(bs))
    check(ccall((:dincae_points_assemble, lib), Int32,
                (Ptr{Cvoid}, Ptr{Int64}, Int32, Int32, UInt64, UInt64, Ptr{UInt8}, Ptr{Int64}, Ptr{Float32}, Int32),
                handle(d), idx, length(bs), d.train, rand(UInt64), UInt64(index), keep, koff, xin, HOST))
    xtrue = [(pos = d.pos[d.perm[j]], x = Atype(d.x[d.perm[j]])) for j in bs]
    return ((Atype(xin), xtrue), bs[end])
end

end # module
