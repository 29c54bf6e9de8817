# DCEGMCuda.jl — Julia host shim over libdcegm_cuda.so (include/dcegm.h).
#
# WRITTEN, NOT EXECUTED: there is no `julia` binary in the build image or on the GPU box (SURVEY.md §8c).
# The same C entry points are exercised from Python/ctypes (dcegm.jl_b200/_lib.py) by tests/.
#
# Usage inside the reference package: `include("DCEGMCuda.jl")` after src/DCEGM.jl's own includes, then
# `DCEGMCuda.dc_EGM!(m, p)` replaces DCEGM.dc_EGM! (src/dc_algo.jl:8-206); `runf`, `rung`, `runfdemo`, `bm`
# keep their bodies (src/dc_algo.jl:335-362, src/bench.jl:3-31).  Types `Param`, `Point`, `MLine`,
# `Envelope`, `FModel`, `GModel`, `Simulation` are the reference's own (src/param.jl, line.jl, envelope.jl,
# simulate.jl) — this file only moves data.
module DCEGMCuda

using ..DCEGM: Param, Point, MLine, Envelope, FModel, GModel, Simulation

const LIB = get(ENV, "DCEGM_CUDA_LIB", "libdcegm_cuda")

# ---- POD mirror of dcegm_param (include/dcegm.h): 23 Float64, 8 Int32, 1 Int64 -------------------------
struct CParam
    gamma::Float64; beta::Float64; sigma::Float64; lambda::Float64; R::Float64
    a_high::Float64; a_low::Float64; a_lowT::Float64; cfloor::Float64; cfloor_plot::Float64
    alpha::Float64; alphaT::Float64; inc0::Float64; inc1::Float64; inc2::Float64
    rho::Float64; pension::Float64; delta::Float64; maxinterest::Float64; bbar::Float64; nu::Float64
    initw0::Float64; initw1::Float64
    na::Int32; ny::Int32; nT::Int32; nD::Int32; k::Int32; retage::Int32; nsims::Int32; quad_group::Int32
    rseed::Int64
end
CParam(p::Param) = CParam(p.gamma, p.beta, p.sigma, p.lambda, p.R, p.a_high, p.a_low, p.a_lowT, p.cfloor,
    p.cfloor_plot, p.alpha, p.alphaT, p.inc0, p.inc1, p.inc2, p.ρ, p.pension, p.delta, p.maxinterest,
    p.bbar, p.ν, p.initw0, p.initw1, p.na, p.ny, p.nT, p.nD, p.k, p.retage, p.nsims, 1, p.rseed)

struct CLayout   # dcegm_layout
    model::Int32; nmodels::Int32; na::Int32; ny::Int32; nT::Int32; nstate::Int32
    cap::Int32; isec_cap::Int32; rem_words::Int32; seg_cap::Int32; has_dirty::Int32; has_lines::Int32
end

struct CSimShocks   # dcegm_sim_shocks
    wshocks::Ptr{Float64}; dshocks::Ptr{Float64}; rshocks::Ptr{Float64}; w0shocks::Ptr{Float64}
    ystate0::Ptr{Int32}; nsims::Int32; _pad::Int32
end

# ---- handle: one per GPU, owns device memory and streams -------------------------------------------------
const HANDLE = Ref{Ptr{Cvoid}}(C_NULL)
function handle(device::Integer = 0)
    if HANDLE[] == C_NULL
        rc = ccall((:dcegm_create, LIB), Cint, (Cint, Ref{Ptr{Cvoid}}), device, HANDLE)
        rc == 0 || error("dcegm_create failed ($rc): no CUDA device — this backend has no CPU path")
        atexit(() -> ccall((:dcegm_destroy, LIB), Cint, (Ptr{Cvoid},), HANDLE[]))
    end
    HANDLE[]
end
lasterr() = unsafe_string(ccall((:dcegm_last_error, LIB), Cstring, (Ptr{Cvoid},), handle()))
check(rc) = rc == 0 || error("libdcegm_cuda ($rc): " * lasterr())

# per-model status bits -> the exceptions the reference would raise (SURVEY.md §5; bit values: include/dcegm.h).
# Informational bits (DCEGM_ST_EXTRA_RIGHT 0x1000, DCEGM_ST_NONMONO_FIRST 0x2000) do not raise: runf returns normally.
const ST_INFO_MASK = 0x1000 | 0x2000
function raise_for_status(st::Integer)
    st & ~ST_INFO_MASK == 0 && return
    st & 0x0004 != 0 && error("knot-vectors must be unique and sorted in increasing order")  # line.jl:236 (decreasing knots)
    st & 0x4000 != 0 && error("knot-vectors must be unique and sorted in increasing order")  # line.jl:236 (repeated knot)
    st & 0x0008 != 0 && throw(ArgumentError("The interval [a,b] is not a bracketing interval"))  # envelope.jl:199
    st & 0x0010 != 0 && throw(BoundsError())                                                     # dc_algo.jl:486-491
    st & 0x0020 != 0 && error("an upper envelope requires by definition at least 2 lines.")  # envelope.jl:149
    st & 0x0400 != 0 && error("neg cons")                                                     # simulate.jl:111-116
    error("dcegm status 0x$(string(st, base = 16))")
end

layout() = (l = Ref{CLayout}(); check(ccall((:dcegm_get_layout, LIB), Cint, (Ptr{Cvoid}, Ref{CLayout}), handle(), l)); l[])

# ---- results: everything of one model in five bulk calls ----------------------------------------------------
# Vector{Point{Float64}} is bit-compatible with the library's interleaved (x,y) doubles.
function fetch_model(model::Integer)
    l = layout()
    nslot = l.nT * l.nstate * 4
    lines = Vector{Point{Float64}}(undef, nslot * l.cap)
    lens = Vector{Int32}(undef, nslot)
    vb = Vector{Float64}(undef, l.nT * l.nstate * 2)
    isec = Vector{Point{Float64}}(undef, l.nT * l.nstate * l.isec_cap)
    meta = Vector{NTuple{4,Int32}}(undef, l.nT * l.nstate)
    rem = Vector{UInt32}(undef, l.nT * l.nstate * l.rem_words)
    h = handle()
    check(ccall((:dcegm_fetch_lines, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Cvoid}), h, model, 1, lines))
    check(ccall((:dcegm_fetch_lens, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Int32}), h, model, 1, lens))
    check(ccall((:dcegm_fetch_vbound, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Float64}), h, model, 1, vb))
    check(ccall((:dcegm_fetch_isects, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Cvoid}), h, model, 1, isec))
    check(ccall((:dcegm_fetch_meta, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Cvoid}), h, model, 1, meta))
    check(ccall((:dcegm_fetch_removed, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{UInt32}), h, model, 1, rem))
    (l, lines, lens, vb, isec, meta, rem)
end

# slot of (it, iy, id, kind), all 1-based except kind (0 = c, 1 = v) — layout comment in include/dcegm.h
slot(l, it, iy, id, kind) = ((((it - 1) * l.nstate + (iy - 1)) * 2 + (id - 1)) * 2 + kind)

function envelope(res, it, iy, id, kind)
    l, lines, lens, vb, isec, meta, rem = res
    s = slot(l, it, iy, id, kind)
    e = Envelope(MLine(lines[s * l.cap + 1 : s * l.cap + lens[s + 1]]))
    if kind == 1
        e.vbound = vb[((it - 1) * l.nstate + (iy - 1)) * 2 + id]
        if id == 1   # worker: intersections and removed indices of secondary_envelope (envelope.jl:268-285)
            st = (it - 1) * l.nstate + (iy - 1)
            e.isects = isec[st * l.isec_cap + 1 : st * l.isec_cap + meta[st + 1][2]]
            w = rem[st * l.rem_words + 1 : (st + 1) * l.rem_words]
            e.removed = [k for k in 1:(32 * l.rem_words) if (w[(k - 1) ÷ 32 + 1] >> ((k - 1) % 32)) & 1 == 1]
        end
    end
    e
end

# ---- dc_EGM!: the drop-in for src/dc_algo.jl:8-105 (FModel) and :112-206 (GModel) ---------------------------
function dc_EGM!(m::FModel, p::Param; demo = false)
    check(ccall((:dcegm_solve_f, LIB), Cint, (Ptr{Cvoid}, Ref{CParam}, Cint, Cuint), handle(), CParam(p), 1, demo ? 1 : 0))
    st = Ref{Int32}(0); check(ccall((:dcegm_status, LIB), Cint, (Ptr{Cvoid}, Ref{Int32}), handle(), st)); raise_for_status(st[])
    res = fetch_model(0)
    for it in 1:p.nT, id in 1:p.nD
        m.c[id, it] = envelope(res, it, 1, id, 0)
        m.v[id, it] = envelope(res, it, 1, id, 1)
    end
    if demo   # vdirty / cdirty (dc_algo.jl:79-82): dcegm_fetch_line with kind 2 / 3
        buf = Vector{Point{Float64}}(undef, p.na); n = Ref{Cint}(0)
        for it in 1:p.nT-1, id in 1:p.nD, (kind, dst) in ((2, m.cdirty), (3, m.vdirty))
            check(ccall((:dcegm_fetch_line, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Cint, Cint, Cint, Ptr{Cvoid}, Cint, Ref{Cint}),
                        handle(), 0, kind, id - 1, 0, it - 1, buf, p.na, n))
            dst[id, it] = MLine(buf[1:n[]])
        end
    end
    m
end

function dc_EGM!(m::GModel, p::Param)
    check(ccall((:dcegm_solve_g, LIB), Cint, (Ptr{Cvoid}, Ref{CParam}, Cint, Cuint), handle(), CParam(p), 1, 0))
    st = Ref{Int32}(0); check(ccall((:dcegm_status, LIB), Cint, (Ptr{Cvoid}, Ref{Int32}), handle(), st)); raise_for_status(st[])
    res = fetch_model(0)
    for it in 1:p.nT, iy in 1:p.ny, id in 1:p.nD
        m.c[id, iy, it] = envelope(res, it, iy, id, 0)
        m.v[id, iy, it] = envelope(res, it, iy, id, 1)
    end
    m
end

# ---- minimal_EGM family (src/minimal_EGM.jl): returns the data the reference plots ---------------------------
function minimal(p::Param, sym::Symbol)
    check(ccall((sym, LIB), Cint, (Ptr{Cvoid}, Ref{CParam}, Cint, Cuint), handle(), CParam(p), 1, 0))
    res = fetch_model(0)
    l, lines, lens = res[1], res[2], res[3]
    line(it, kind) = (s = slot(l, it, 1, 1, kind); lines[s * l.cap + 1 : s * l.cap + lens[s + 1]])
    m = [[q.x for q in line(it, 0)] for it in 1:p.nT]
    c = [[q.y for q in line(it, 0)] for it in 1:p.nT]
    v = sym == :dcegm_solve_minimal_bequest ? [[q.y for q in line(it, 1)] for it in 1:p.nT] : nothing
    m, c, v
end
minimal_EGM(p::Param) = minimal(p, :dcegm_solve_minimal)[1:2]
minimal_EGM_borrow(p::Param) = minimal(p, :dcegm_solve_minimal_borrow)[1:2]
minimal_EGM_bequest(p::Param) = minimal(p, :dcegm_solve_minimal_bequest)

# ---- sim (src/simulate.jl:61-203) on the solution resident after dc_EGM! ------------------------------------
# shocks = nothing: counter-based Philox draws keyed by p.rseed (csrc/dc_rng.h).  To follow a Julia RNG
# stream pass the raw rand() matrices, e.g. (Random.seed!(p.rseed); w = rand(p.nsims,p.nT); d = rand(...); ...).
function sim(m::Union{FModel,GModel}, p::Param; shocks = nothing)
    sh = shocks === nothing ? C_NULL :
         Ref(CSimShocks(pointer(shocks.w), pointer(shocks.d), m isa FModel ? pointer(shocks.r) : C_NULL, pointer(shocks.w0),
                        m isa GModel ? pointer(shocks.y0) : C_NULL, p.nsims, 0))
    GC.@preserve shocks check(ccall((:dcegm_simulate, LIB), Cint, (Ptr{Cvoid}, Cuint, Ptr{Cvoid}), handle(), 1, sh))
    s = Simulation(p)
    ys = Matrix{Int32}(undef, p.nsims, p.nT); wk = similar(ys); ra = Vector{Int32}(undef, p.nsims)
    check(ccall((:dcegm_fetch_sim_panel, LIB), Cint,
        (Ptr{Cvoid}, Cint, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}, Ptr{Int32}, Ptr{Int32}),
        handle(), 0, s.w0, s.w1, s.cons, s.util, s.inc, s.prob_work, ys, wk, ra))
    s.ystate .= ys; s.worker .= wk; s.ret_age .= ra
    st = Ref{Int32}(0)
    check(ccall((:dcegm_fetch_sim_moments, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Float64}, Ptr{Int32}, Ref{Int32}), handle(), 0, 1, C_NULL, C_NULL, st))
    raise_for_status(st[])
    s
end

# ---- parameter sweeps: a Vector{Param} in one call, all lines packed into one pinned host buffer -------------
function solve_batch(ps::Vector{Param}; model = :f)
    cp = CParam.(ps); l0 = ps[1]
    nstate = model == :g ? l0.ny : 1
    nslot = length(ps) * l0.nT * nstate * 4
    cap = 8 * cld(l0.na + l0.na ÷ 10 + 2 * max(l0.na ÷ 16, 32) + 2, 8)
    xy = Vector{Point{Float64}}(undef, nslot * cap); off = Vector{Int64}(undef, nslot + 1); np = Ref{Int64}(0)
    f = model == :g ? :dcegm_solve_g_packed : :dcegm_solve_f_packed
    check(ccall((f, LIB), Cint, (Ptr{Cvoid}, Ptr{CParam}, Cint, Cuint, Ptr{Cvoid}, Csize_t, Ptr{Int64}, Ref{Int64}),
                handle(), cp, length(cp), 0, xy, length(xy), off, np))
    st = Vector{Int32}(undef, length(ps)); check(ccall((:dcegm_status, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), handle(), st))
    resize!(xy, np[]); xy, off, st   # line s (0-based slot order of include/dcegm.h) = xy[off[s+1]+1 : off[s+2]]
end

# ---- the same sweep in the compact line format (include/dcegm.h "compact lines"): half the bytes over PCIe; lines are
# rebuilt lazily, one Envelope pair at a time, by the handle-free decoder.  woff[slot] is the START of a slot's record;
# lines shared bit for bit with the chunk's first model (retirees in a (lambda, sigma) sweep) point at its records ----
struct CompactBatch
    words::Vector{Float64}; woff::Vector{Int64}; lens::Vector{Int32}; aux::Vector{Int32}
    grids::Vector{Vector{Float64}}   # savings grid of every model (dcegm_get_grid)
    nT::Int; nstate::Int
end

function solve_batch_compact(ps::Vector{Param}; model = :f)
    cp = CParam.(ps); l0 = ps[1]
    nstate = model == :g ? l0.ny : 1
    nslot = length(ps) * l0.nT * nstate * 4
    cap = 8 * cld(l0.na + l0.na ÷ 10 + 2 * max(l0.na ÷ 16, 32) + 2, 8)
    words = Vector{Float64}(undef, nslot * (2 * cap + cap ÷ 64 + 8)); woff = Vector{Int64}(undef, nslot + 1)
    lens = Vector{Int32}(undef, nslot); aux = Vector{Int32}(undef, nslot); nw = Ref{Int64}(0)
    f = model == :g ? :dcegm_solve_g_compact : :dcegm_solve_f_compact
    check(ccall((f, LIB), Cint, (Ptr{Cvoid}, Ptr{CParam}, Cint, Cuint, Ptr{Float64}, Csize_t, Ptr{Int64}, Ptr{Int32}, Ptr{Int32}, Ref{Int64}),
                handle(), cp, length(cp), 0, words, length(words), woff, lens, aux, nw))
    resize!(words, nw[])
    grids = map(0:length(ps)-1) do m
        na = Ref{Cint}(0); g = Vector{Float64}(undef, ps[m + 1].na)
        check(ccall((:dcegm_get_grid, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Float64}, Cint, Ref{Cint}), handle(), m, g, length(g), na)); g
    end
    CompactBatch(words, woff, lens, aux, grids, l0.nT, nstate)
end

# policy and value line (Vector{Point}) of model m (1-based), period it, state iy, choice id
function lines(b::CompactBatch, m, it, iy, id)
    sc = ((((m - 1) * b.nT + (it - 1)) * b.nstate + (iy - 1)) * 2 + (id - 1)) * 2   # 0-based slot of the policy line
    nc, nv = b.lens[sc + 1], b.lens[sc + 2]
    c = Vector{Point{Float64}}(undef, nc); v = Vector{Point{Float64}}(undef, nv); g = b.grids[m]
    rc = ccall((:dcegm_decode_compact_pair, LIB), Cint,
               (Ptr{Float64}, Cint, Ptr{Float64}, Cint, Cint, Ptr{Float64}, Cint, Cint, Ptr{Cvoid}, Ptr{Cvoid}),
               g, length(g), pointer(b.words, b.woff[sc + 1] + 1), nc, b.aux[sc + 1], pointer(b.words, b.woff[sc + 2] + 1), nv, b.aux[sc + 2], c, v)
    rc == 0 || error("inconsistent compact record")
    MLine(c), MLine(v)
end

end # module
