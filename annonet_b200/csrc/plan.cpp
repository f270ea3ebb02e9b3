#include "plan.h"

#include <algorithm>
#include <cmath>
#include <map>

namespace ab {

namespace {

struct Atom { int level; int channels; bool dead; int c0; };
struct Seg { int atom; };
struct Sym { std::vector<int> atoms; int level = 0; bool pooled = false; bool is_image = false; };

}  // namespace

LevelDims level_dims(int h, int w, int levels)
{
    LevelDims d{};
    d.h[0] = h; d.w[0] = w;
    for (int n = 1; n <= levels; ++n) {
        if (d.h[n - 1] < 3 || d.w[n - 1] < 3) throw std::invalid_argument("input smaller than the required input dimension");
        d.h[n] = 1 + (d.h[n - 1] - 3) / 2;
        d.w[n] = 1 + (d.w[n - 1] - 3) / 2;
    }
    d.first_resized = -1;
    for (int n = levels - 1; n >= 0; --n) {
        d.up_h[n] = 2 * d.h[n + 1] + 1;
        d.up_w[n] = 2 * d.w[n + 1] + 1;
        d.resized[n] = d.up_h[n] != d.h[n] || d.up_w[n] != d.w[n];   // (by construction the difference is 0 or -1 per axis)
        if (d.resized[n]) d.first_resized = n;
    }
    d.up_h[levels] = d.h[levels]; d.up_w[levels] = d.w[levels]; d.resized[levels] = false;
    return d;
}

ResizeAxis resize_axis(int n_in, int n_out, bool simd_columns)
{
    ResizeAxis a;
    a.lo.assign(n_out, 0); a.hi.assign(n_out, 0); a.frac.assign(n_out, 0.f); a.ifrac.assign(n_out, 1.f);
    const double scale = (n_in - 1) / (double)std::max(n_out - 1, 1);
    int c = 0;
    if (simd_columns) {
        const double x0 = -4 * scale;
        volatile float xv[4] = {(float)x0, (float)(x0 + scale), (float)(x0 + 2 * scale), (float)(x0 + 3 * scale)};
        const float step = (float)(4 * scale);
        while (true) {
            int left[4];
            for (int j = 0; j < 4; ++j) { xv[j] = xv[j] + step; left[j] = (int)xv[j]; }
            if (left[3] + 1 >= n_in || c + 4 > n_out) break;
            for (int j = 0; j < 4; ++j) {
                a.lo[c + j] = left[j]; a.hi[c + j] = left[j] + 1;
                volatile float f = xv[j] - (float)left[j];
                a.frac[c + j] = f; a.ifrac[c + j] = 1.f - f;
            }
            c += 4;
        }
    }
    double x = -scale + c * scale;
    for (; c < n_out; ++c) {
        x += scale;
        const int left = (int)std::floor(x);
        a.lo[c] = left; a.hi[c] = std::min(left + 1, n_in - 1);
        a.frac[c] = (float)(x - left);
        // rows: simd4f(1 - tb_frac) rounds the DOUBLE difference; tail columns: float lr_frac, then 1 - lr_frac
        a.ifrac[c] = simd_columns ? 1.f - a.frac[c] : (float)(1.0 - (x - left));
    }
    return a;
}

Plan build_plan(const NetModel& net)
{
    const std::vector<Layer>& L = net.layers;
    const int n = (int)L.size();
    Plan plan;
    plan.levels = net.levels;
    plan.class_count = net.class_count();
    plan.layer_off.assign(n, 0);
    plan.layer_size.assign(n, 0);

    std::vector<Atom> atoms;
    std::vector<Sym> sym(n);
    std::vector<int> producer;           // atom -> group index
    std::vector<std::vector<int>> logical_in;  // group -> atoms (logical order)
    sym[n - 1].is_image = true;
    sym[n - 1].level = 0;

    auto resolve = [&](int i) { while (L[i].type == L_TAG) ++i; return i; };

    for (int i = n - 2; i >= 1; --i) {
        const Layer& l = L[i];
        switch (l.type) {
        case L_TAG:
        case L_MISH:
        case L_BN:
        case L_AFFINE: {
            sym[i] = sym[resolve(i + 1)];
            if (l.type == L_TAG) break;
            if (sym[i].atoms.size() != 1) throw std::logic_error("mish/bn on a multi-tensor input");
            ConvGroup& g = plan.groups[producer[sym[i].atoms[0]]];
            if (l.type == L_MISH) {
                if (g.has_mish) throw std::logic_error("two mish layers on one conv");
                g.has_mish = true;
                g.bn_first = g.dlib_bn >= 0;
            } else {
                if (g.dlib_bn >= 0) throw std::logic_error("two bn layers on one conv");
                g.dlib_bn = i;
            }
            break;
        }
        case L_MAXPOOL: {
            sym[i] = sym[resolve(i + 1)];
            if (sym[i].pooled) throw std::logic_error("pool of a pool");
            sym[i].pooled = true;
            sym[i].level += 1;
            break;
        }
        case L_CONCAT_PREV: {
            const Sym& a = sym[resolve(i + 1)];
            const Sym& b = sym[resolve(find_tag(L, i, l.tag_id))];
            if (a.level != b.level || a.pooled || b.pooled) throw std::logic_error("concat_prev of different resolutions");
            sym[i] = a;
            sym[i].atoms.insert(sym[i].atoms.end(), b.atoms.begin(), b.atoms.end());
            break;
        }
        case L_CONCAT: {
            sym[i] = Sym();
            bool first = true;
            for (int t : l.tags) {
                const Sym& b = sym[resolve(find_tag(L, i, t))];
                if (first) { sym[i].level = b.level; first = false; }
                else if (b.level != sym[i].level) throw std::logic_error("concat of different resolutions");
                sym[i].atoms.insert(sym[i].atoms.end(), b.atoms.begin(), b.atoms.end());
            }
            break;
        }
        case L_CON:
        case L_CONT: {
            const Sym& in = sym[resolve(i + 1)];
            ConvGroup g;
            g.dlib_conv = i;
            g.is_cont = l.type == L_CONT;
            g.ksize = (int)l.nr; g.stride = l.stride; g.pad = l.pad;
            g.pool_input = in.pooled;
            g.dead = l.dead;
            g.cout = (int)l.num_filters;
            int out_level = in.level;
            if (g.is_cont && g.stride == 2) out_level = in.level - 1;
            if (in.is_image) { g.in_level = -1; g.cin = 3; }
            else {
                g.in_level = in.pooled ? in.level - 1 : in.level;
                for (int a : in.atoms) g.cin += atoms[a].channels;
            }
            g.out_level = (i == 1) ? -2 : out_level;
            Atom at{out_level, g.cout, l.dead, 0};
            atoms.push_back(at);
            producer.push_back((int)plan.groups.size());
            logical_in.push_back(in.atoms);
            plan.groups.push_back(g);
            sym[i] = Sym();
            sym[i].atoms = {(int)atoms.size() - 1};
            sym[i].level = out_level;
            break;
        }
        default: throw std::logic_error("unexpected layer type");
        }
    }

    // physical channel offsets: per level, live atoms newest first; the logits atom lives elsewhere
    plan.level_channels.assign(plan.levels + 1, 0);
    for (int a = (int)atoms.size() - 1; a >= 0; --a) {
        const ConvGroup& g = plan.groups[producer[a]];
        if (g.out_level == -2) continue;
        if (atoms[a].dead) { plan.max_dead_channels = std::max(plan.max_dead_channels, atoms[a].channels); continue; }
        atoms[a].c0 = plan.level_channels[atoms[a].level];
        plan.level_channels[atoms[a].level] += atoms[a].channels;
    }

    // input slices + permutations, parameter offsets
    size_t off = 0;
    for (size_t gi = 0; gi < plan.groups.size(); ++gi) {
        ConvGroup& g = plan.groups[gi];
        const int a_out = (int)gi;  // one atom per group, same order
        g.out_c0 = atoms[a_out].dead || g.out_level == -2 ? 0 : atoms[a_out].c0;
        if (g.in_level >= 0) {
            std::vector<int> phys;
            for (int a : logical_in[gi]) {
                if (atoms[a].dead) throw std::logic_error("a dead tensor is consumed");
                for (int c = 0; c < atoms[a].channels; ++c) phys.push_back(atoms[a].c0 + c);
            }
            const int lo = *std::min_element(phys.begin(), phys.end());
            const int hi = *std::max_element(phys.begin(), phys.end());
            if (hi - lo + 1 != (int)phys.size()) throw std::logic_error("conv input is not a contiguous channel range");
            g.in_c0 = lo;
            g.perm.resize(phys.size());
            for (size_t c = 0; c < phys.size(); ++c) g.perm[c] = phys[c] - lo;
        } else {
            g.in_c0 = 0;
            g.perm = {0, 1, 2};
        }
        g.w_off = off;
        g.b_off = off + g.nweights();
        plan.layer_off[g.dlib_conv] = off;
        plan.layer_size[g.dlib_conv] = g.nweights() + g.cout;
        off += g.nweights() + g.cout;
        if (g.dlib_bn >= 0) {
            g.bn_off = off;
            plan.layer_off[g.dlib_bn] = off;
            plan.layer_size[g.dlib_bn] = 2 * (size_t)g.cout;
            off += 2 * (size_t)g.cout;
        }
    }
    plan.num_params = off;
    return plan;
}

void pack_filters(const ConvGroup& g, const float* p, float* w, float* b)
{
    const int kk = g.ksize * g.ksize;
    if (!g.is_cont) {  // dlib [cout][cin][k][k]
        for (int co = 0; co < g.cout; ++co)
            for (int ci = 0; ci < g.cin; ++ci)
                for (int t = 0; t < kk; ++t) w[((size_t)co * g.cin + g.perm[ci]) * kk + t] = p[((size_t)co * g.cin + ci) * kk + t];
    } else {           // dlib [cin][cout][k][k]
        for (int ci = 0; ci < g.cin; ++ci)
            for (int co = 0; co < g.cout; ++co)
                for (int t = 0; t < kk; ++t) w[((size_t)g.perm[ci] * g.cout + co) * kk + t] = p[((size_t)ci * g.cout + co) * kk + t];
    }
    for (int co = 0; co < g.cout; ++co) b[co] = p[g.nweights() + co];
}

void unpack_filters(const ConvGroup& g, const float* w, const float* b, float* p)
{
    const int kk = g.ksize * g.ksize;
    if (!g.is_cont) {
        for (int co = 0; co < g.cout; ++co)
            for (int ci = 0; ci < g.cin; ++ci)
                for (int t = 0; t < kk; ++t) p[((size_t)co * g.cin + ci) * kk + t] = w[((size_t)co * g.cin + g.perm[ci]) * kk + t];
    } else {
        for (int ci = 0; ci < g.cin; ++ci)
            for (int co = 0; co < g.cout; ++co)
                for (int t = 0; t < kk; ++t) p[((size_t)ci * g.cout + co) * kk + t] = w[((size_t)g.perm[ci] * g.cout + co) * kk + t];
    }
    for (int co = 0; co < g.cout; ++co) p[g.nweights() + co] = b[co];
}

}  // namespace ab
