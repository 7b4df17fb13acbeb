// fast_bvh.cpp -- binned-SAH binary build over the scene's leaf objects, collapsed to 4-wide.
// Host-side set-up code (like the reference's BvhNode constructor, bvhNode.h:25-59); nothing here
// runs per ray.  See fast_bvh.h for what the structure is allowed to decide (culling only).
#include "fast_bvh.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace rtow {
namespace {

struct DBox {
    double lo[3], hi[3];
    DBox() { for (int a = 0; a < 3; ++a) { lo[a] = 1e300; hi[a] = -1e300; } }
    void grow(const DBox &b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void grow(const double p[3]) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    bool valid() const { return lo[0] <= hi[0]; }
    double area() const {
        double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return 2 * (dx * dy + dy * dz + dz * dx);
    }
};

// Box of `inner` (given in the space below the chain) expressed in the space above the chain.
// chain ops map the RAY outer->inner (hitable.h:53-62,125-154): Translate o' = o - offset,
// RotateY o' = (c*x - s*z, y, s*x + c*z).  Points go back with the inverse maps, innermost first.
DBox chain_box(const rtow_xform *xf, int n, DBox b) {
    for (int i = n - 1; i >= 0; --i) {
        if (xf[i].type == RTOW_XF_TRANSLATE) {
            for (int a = 0; a < 3; ++a) { b.lo[a] += xf[i].f[a]; b.hi[a] += xf[i].f[a]; }
        } else if (xf[i].type == RTOW_XF_ROTATE_Y) {
            double s = xf[i].f[0], c = xf[i].f[1];
            DBox o;
            for (int k = 0; k < 8; ++k) {
                double x = (k & 1) ? b.hi[0] : b.lo[0], y = (k & 2) ? b.hi[1] : b.lo[1], z = (k & 4) ? b.hi[2] : b.lo[2];
                double p[3] = {c * x + s * z, y, -s * x + c * z};
                o.grow(p);
            }
            b = o;
        }
    }
    return b;
}

bool prim_box(const rtow_scene_desc *d, int pid, DBox &out, float *tlo, float *thi, int depth = 0) {
    const rtow_prim &p = d->prims[pid];
    DBox b;
    switch (p.type) {
    case RTOW_PRIM_SPHERE:
        for (int a = 0; a < 3; ++a) { b.lo[a] = (double)p.f[a] - std::fabs((double)p.f[3]); b.hi[a] = (double)p.f[a] + std::fabs((double)p.f[3]); }
        break;
    case RTOW_PRIM_MOVING_SPHERE: {
        double r = std::fabs((double)p.f[8]);
        for (int a = 0; a < 3; ++a) {
            b.lo[a] = std::min((double)p.f[a], (double)p.f[3 + a]) - r;
            b.hi[a] = std::max((double)p.f[a], (double)p.f[3 + a]) + r;
        }
        // centre(time) stays inside this box only for time in [time0, time1] (sphere.h:129-131)
        float a0 = std::min(p.f[6], p.f[7]), a1 = std::max(p.f[6], p.f[7]);
        *tlo = std::max(*tlo, a0); *thi = std::min(*thi, a1);
        break;
    }
    case RTOW_PRIM_RECT_XY: b.lo[0] = p.f[0]; b.hi[0] = p.f[1]; b.lo[1] = p.f[2]; b.hi[1] = p.f[3]; b.lo[2] = b.hi[2] = p.f[4]; break;
    case RTOW_PRIM_RECT_XZ: b.lo[0] = p.f[0]; b.hi[0] = p.f[1]; b.lo[2] = p.f[2]; b.hi[2] = p.f[3]; b.lo[1] = b.hi[1] = p.f[4]; break;
    case RTOW_PRIM_RECT_YZ: b.lo[1] = p.f[0]; b.hi[1] = p.f[1]; b.lo[2] = p.f[2]; b.hi[2] = p.f[3]; b.lo[0] = b.hi[0] = p.f[4]; break;
    case RTOW_PRIM_MEDIUM: {
        if (depth > 0) return false;
        for (int c = p.child_begin; c < p.child_begin + p.child_count; ++c) {
            DBox cb;
            if (!prim_box(d, c, cb, tlo, thi, depth + 1)) return false;
            b.grow(cb);
        }
        break;
    }
    default: return false;
    }
    for (int a = 0; a < 3; ++a) {
        if (!(b.lo[a] <= b.hi[a]) || !std::isfinite(b.lo[a]) || !std::isfinite(b.hi[a])) return false;
    }
    out = chain_box(d->xforms + p.xform_begin, p.xform_count, b);
    return true;
}

struct BNode {
    DBox box;
    int left = -1, right = -1;   // binary node indices
    int obj = -1;                // >= 0: leaf
    int cnt = 0;                 // objects below
};

// Runs f(i) for i in [0, n) on up to `threads` host threads (indices are claimed from a counter; f must not care who runs it).
template <class F>
void parallel_for(int n, int threads, const F &f) {
    threads = std::max(1, std::min(threads, n));
    if (threads == 1) { for (int i = 0; i < n; ++i) f(i); return; }
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    auto work = [&] { for (int i = next++; i < n; i = next++) f(i); };
    try {
        for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    } catch (...) {
        // the system would not start another thread: the counter hands its share to the ones that run
    }
    work();
    for (auto &t : pool) t.join();
}
int host_threads() {
    static const int n = [] {
        const char *e = getenv("RTOW_BVH_THREADS");
        if (e && atoi(e) > 0) return atoi(e);
        const unsigned hc = std::thread::hardware_concurrency();
        return (int)std::max(1u, std::min(hc ? hc : 1u, 32u));
    }();
    return n;
}

struct Builder {
    const std::vector<DBox> &boxes;
    std::vector<double> cx, cy, cz;
    std::vector<int> idx;
    std::vector<BNode> nodes;
    // Subtrees of at most `cutoff` objects are built as independent tasks on the host's threads (the top of the tree, whose
    // splits sweep the whole object array, is built first).  A subtree over m objects has exactly 2m - 1 nodes, so every
    // task writes its nodes straight into its own block of the one node array.  The split of the work depends on the
    // object count alone, so the tree -- and its node numbering -- is the same on every machine and every run.
    struct Task { int b, e, parent, side, base; };
    std::vector<Task> tasks;
    int cutoff = 0;

    explicit Builder(const std::vector<DBox> &b) : boxes(b) {
        size_t n = b.size();
        cx.resize(n); cy.resize(n); cz.resize(n); idx.resize(n);
        for (size_t i = 0; i < n; ++i) {
            cx[i] = 0.5 * (b[i].lo[0] + b[i].hi[0]); cy[i] = 0.5 * (b[i].lo[1] + b[i].hi[1]); cz[i] = 0.5 * (b[i].lo[2] + b[i].hi[2]);
            idx[i] = (int)i;
        }
    }
    double cen(int o, int a) const { return a == 0 ? cx[o] : a == 1 ? cy[o] : cz[o]; }

    int build_all() {
        const int n = (int)boxes.size();
        nodes.assign((size_t)2 * n - 1, BNode());
        int cursor = 0;
        if (n < (1 << 15)) return build(0, n, cursor, -1, 0);
        cutoff = std::max(1 << 12, n >> 7);
        std::vector<BNode> all;
        all.swap(nodes);                       // the top of the tree goes to a small array of its own first
        nodes.assign(4 * (size_t)(n / cutoff + 1) + 8, BNode());
        const int root = build(0, n, cursor, -1, 0);
        const int n_top = cursor;
        std::vector<BNode> top;
        top.swap(nodes);
        nodes.swap(all);
        int next = n_top;
        for (Task &k : tasks) { k.base = next; next += 2 * (k.e - k.b) - 1; }
        for (int i = 0; i < n_top; ++i) nodes[i] = top[i];
        for (const Task &k : tasks) (k.side ? nodes[k.parent].right : nodes[k.parent].left) = k.base;
        parallel_for((int)tasks.size(), host_threads(), [&](int t) {
            int cur = tasks[t].base;
            build(tasks[t].b, tasks[t].e, cur, -2, 0);
        });
        return root;
    }

    // parent >= 0 (or -1 for the root): building the shared top of the tree, ranges of <= cutoff objects become tasks;
    // parent == -2: inside a task (no further tasks)
    int build(int b, int e, int &cursor, int parent, int side) {
        std::vector<BNode> &out = nodes;
        if (parent != -2 && cutoff > 0 && parent >= 0 && e - b <= cutoff) {
            tasks.push_back(Task{b, e, parent, side, -1});
            return -1;      // patched when the task's nodes are appended
        }
        const int me = cursor++;
        // (the top of the tree is built into an array sized for balanced splits; a scene whose splits peel a few objects off
        // per level needs more -- only this serial phase may grow it: a task's block is exactly 2m - 1 nodes)
        if (parent != -2 && (size_t)me >= out.size()) out.resize(std::max(out.size() * 2, (size_t)me + 1));
        // the two sweeps over the range (bounds, then bins) are reductions of exact min / max / counts: on the big ranges of
        // the serial top phase they run in slices on the host's threads and give the same numbers
        const int T = (parent != -2 && cutoff > 0 && e - b > cutoff) ? host_threads() : 1;
        auto slice = [&](int t, int &lo, int &hi) { lo = b + (int)((long long)(e - b) * t / T); hi = b + (int)((long long)(e - b) * (t + 1) / T); };
        DBox bb, cb;
        {
            std::vector<DBox> pb(T), pc(T);
            parallel_for(T, T, [&](int t) {
                int lo, hi; slice(t, lo, hi);
                DBox x, y;
                for (int i = lo; i < hi; ++i) {
                    x.grow(boxes[idx[i]]);
                    double c[3] = {cx[idx[i]], cy[idx[i]], cz[idx[i]]};
                    y.grow(c);
                }
                pb[t] = x; pc[t] = y;
            });
            for (int t = 0; t < T; ++t) { if (pb[t].valid()) bb.grow(pb[t]); if (pc[t].valid()) cb.grow(pc[t]); }
        }
        out[me].box = bb;
        out[me].cnt = e - b;
        if (e - b == 1) { out[me].obj = idx[b]; return me; }
        const int NB = 16;
        struct Bins { DBox box[3][NB]; int n[3][NB]; };
        bool use[3]; double blo[3], bscale[3];
        for (int a = 0; a < 3; ++a) {
            const double ext = cb.hi[a] - cb.lo[a];
            use[a] = ext > 0; blo[a] = cb.lo[a]; bscale[a] = use[a] ? NB / ext : 0.0;
        }
        std::vector<Bins> part(T);
        parallel_for(T, T, [&](int t) {
            int lo, hi; slice(t, lo, hi);
            Bins &S = part[t];
            for (int a = 0; a < 3; ++a) for (int k = 0; k < NB; ++k) S.n[a][k] = 0;
            for (int i = lo; i < hi; ++i) {
                const int o = idx[i];
                for (int a = 0; a < 3; ++a) {
                    if (!use[a]) continue;
                    int k = (int)((cen(o, a) - blo[a]) * bscale[a]);
                    k = k < 0 ? 0 : k >= NB ? NB - 1 : k;
                    S.n[a][k]++; S.box[a][k].grow(boxes[o]);
                }
            }
        });
        Bins &B0 = part[0];
        for (int t = 1; t < T; ++t)
            for (int a = 0; a < 3; ++a) for (int k = 0; k < NB; ++k) if (part[t].n[a][k]) { B0.n[a][k] += part[t].n[a][k]; B0.box[a][k].grow(part[t].box[a][k]); }
        int best_axis = -1, best_bin = -1;
        double best_cost = 1e300;
        for (int a = 0; a < 3; ++a) {
            if (!use[a]) continue;
            const DBox *bin_box = B0.box[a]; const int *bin_n = B0.n[a];
            double right_area[NB]; int right_n[NB];
            DBox acc; int cnt = 0;
            for (int k = NB - 1; k > 0; --k) {
                if (bin_n[k]) acc.grow(bin_box[k]);
                cnt += bin_n[k];
                right_area[k] = cnt ? acc.area() : 0; right_n[k] = cnt;
            }
            DBox accl; int cl = 0;
            for (int k = 0; k < NB - 1; ++k) {
                if (bin_n[k]) accl.grow(bin_box[k]);
                cl += bin_n[k];
                if (cl == 0 || right_n[k + 1] == 0) continue;
                double cost = accl.area() * cl + right_area[k + 1] * right_n[k + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; }
            }
        }
        int mid;
        if (best_axis < 0) {
            mid = (b + e) / 2;   // all centroids coincide
        } else {
            double lo = cb.lo[best_axis], scale = NB / (cb.hi[best_axis] - cb.lo[best_axis]);
            int a = best_axis, kb = best_bin;
            auto it = std::partition(idx.begin() + b, idx.begin() + e, [&](int o) {
                int k = (int)((cen(o, a) - lo) * scale);
                k = k < 0 ? 0 : k >= NB ? NB - 1 : k;
                return k <= kb;
            });
            mid = (int)(it - idx.begin());
            if (mid == b || mid == e) mid = (b + e) / 2;
        }
        const int next_parent = parent == -2 ? -2 : me;
        int l = build(b, mid, cursor, next_parent, 0);
        int r = build(mid, e, cursor, next_parent, 1);
        out[me].left = l; out[me].right = r;
        return me;
    }
};

inline float i2f(int32_t v) { float f; std::memcpy(&f, &v, 4); return f; }

// outward-rounded float bounds with a little slack for the float arithmetic of the hit code
inline float down(double v, double pad) { float f = (float)(v - pad); while ((double)f > v - pad) f = std::nextafterf(f, -FLT_MAX); return f; }
inline float up(double v, double pad) { float f = (float)(v + pad); while ((double)f < v + pad) f = std::nextafterf(f, FLT_MAX); return f; }

}  // namespace

bool build_fast_bvh(const rtow_scene_desc *d, FastBvh &out, std::string *why) {
    auto fail = [&](const char *m) { if (why) *why = m; return false; };
    if (d->top_kind != RTOW_TOP_BVH || d->n_objects < 1) return fail("not a BVH scene");
    const int n = d->n_objects;
    float tlo = -FLT_MAX, thi = FLT_MAX;

    std::vector<DBox> boxes(n);
    DBox world;
    for (int o = 0; o < n; ++o) {
        const rtow_object &ob = d->objects[o];
        if (ob.prim_count < 1) return fail("empty object");
        for (int p = ob.prim_begin; p < ob.prim_begin + ob.prim_count; ++p) {
            DBox pb;
            if (!prim_box(d, p, pb, &tlo, &thi)) return fail("primitive without a finite box");
            boxes[o].grow(pb);
        }
        world.grow(boxes[o]);
    }
    double ext = 0;
    for (int a = 0; a < 3; ++a) ext = std::max(ext, std::max(std::fabs(world.lo[a]), std::fabs(world.hi[a])));
    for (int a = 0; a < 3; ++a) out.center[a] = (float)(0.5 * (world.lo[a] + world.hi[a]));
    double rad = 0;
    for (int a = 0; a < 3; ++a) { double h = 0.5 * (world.hi[a] - world.lo[a]); rad += h * h; }
    out.radius = (float)(std::sqrt(rad) * 1.0001 + 1e-6);
    out.time_lo = tlo; out.time_hi = thi;

    // guard boxes: the reference node that owns each leaf (bvhNode.h:12-23)
    std::vector<int> owner(n, -1);
    std::vector<char> twice(n, 0);
    for (int i = 0; i < d->n_nodes; ++i) {
        const rtow_bvh_node &nd = d->nodes[i];
        if (nd.left < 0) { if (owner[~nd.left] >= 0) return fail("object referenced by two nodes"); owner[~nd.left] = i; }
        if (nd.right < 0 && nd.right != nd.left) { if (owner[~nd.right] >= 0) return fail("object referenced by two nodes"); owner[~nd.right] = i; }
        if (nd.left < 0 && nd.right == nd.left) twice[~nd.left] = 1;
    }
    if (n >= FAST_LEAF_ID_MASK) return fail("too many objects for the leaf reference encoding");
    out.spheres.assign((size_t)n + 1, make_float4(0, 0, 0, 0));
    out.spheres[n] = make_float4(0, 0, 0, std::nanf(""));   // the never-hit sphere behind unused child slots (NaN discriminant)
    out.guards.assign((size_t)n * 2, make_float4(0, 0, 0, 0));
    std::vector<int32_t> leaf_ref(n);
    for (int o = 0; o < n; ++o) {
        if (owner[o] < 0) return fail("object not referenced by the tree");
        const rtow_bvh_node &g = d->nodes[owner[o]];
        const rtow_object &ob = d->objects[o];
        const rtow_prim &p0 = d->prims[ob.prim_begin];
        bool simple = ob.prim_count == 1 && p0.type == RTOW_PRIM_SPHERE && p0.xform_count == 0;
        bool moving = ob.prim_count == 1 && p0.type == RTOW_PRIM_MOVING_SPHERE && p0.xform_count == 0;
        bool has_medium = false;
        for (int p = ob.prim_begin; p < ob.prim_begin + ob.prim_count; ++p) has_medium |= d->prims[p].type == RTOW_PRIM_MEDIUM;
        int32_t code = o;
        if (simple) code |= FAST_LEAF_SPHERE;
        if (moving) code |= FAST_LEAF_MOVING;
        if (twice[o] && has_medium) code |= FAST_LEAF_TWICE;
        leaf_ref[o] = ~code;
        out.guards[(size_t)o * 2 + 0] = make_float4(g.bmin[0], g.bmin[1], g.bmin[2], 0.f);
        out.guards[(size_t)o * 2 + 1] = make_float4(g.bmax[0], g.bmax[1], g.bmax[2], 0.f);
        if (simple) out.spheres[o] = make_float4(p0.f[0], p0.f[1], p0.f[2], p0.f[3]);
        if (moving) {
            // rtow_prim: f[0..2] centre0, f[3..5] centre1, f[6] time0, f[7] time1, f[8] radius.  centre(time) =
            // c0 + ((time - t0) / (t1 - t0)) * (c1 - c0) (sphere.h:129-131): the two differences do not depend on the ray
            if (out.moving.empty()) out.moving.assign((size_t)n * 2, make_float4(0, 0, 0, 0));
            volatile float dx = p0.f[3] - p0.f[0], dy = p0.f[4] - p0.f[1], dz = p0.f[5] - p0.f[2], dt = p0.f[7] - p0.f[6];   // single float subtractions, as the device does them
            out.spheres[o] = make_float4(p0.f[0], p0.f[1], p0.f[2], p0.f[8]);
            out.moving[(size_t)o * 2 + 0] = make_float4(dx, dy, dz, p0.f[6]);
            out.moving[(size_t)o * 2 + 1] = make_float4(dt, 0.f, 0.f, 0.f);
        }
    }

    Builder B(boxes);
    int broot = B.build_all();

    // collapse to 4-wide, depth first, into wide nodes with float boxes; they are quantised afterwards.
    // A wide node takes its id when it is popped from the stack, so ids are the pre-order of the wide tree with the children
    // taken last slot first, and a subtree occupies one contiguous run of ids.  On big scenes that is used to collapse the
    // subtrees below a frontier on the host's threads: the top is walked serially and only records, in pop order, its own
    // nodes and the frontier subtrees; each of those is collapsed into a block of its own with ids from 0; the run of a
    // block starts at the number of nodes popped before it.  Same ids, same node array as the one-thread walk.
    const double pad_rel = 4e-7;
    struct Wide { float lo[3][4], hi[3][4]; int32_t ref[4]; int nk; };
    struct Item { int bnode; int slot_node; int slot; int depth; };
    auto empty_node = [] {
        Wide w;
        std::memset(&w, 0, sizeof w);
        for (int k = 0; k < 4; ++k) w.ref[k] = FAST_EMPTY;
        return w;
    };
    // fills wide node `w` (id `id` in its array) from binary node it.bnode; internal children go to `stack` in slot order
    auto expand = [&](const Item &it, int id, Wide &w, std::vector<Item> &stack) {
        int kids[4]; int nk = 0;
        const BNode &bn = B.nodes[it.bnode];
        if (bn.obj >= 0) { kids[nk++] = it.bnode; }
        else {
            kids[nk++] = bn.left; kids[nk++] = bn.right;
            while (nk < 4) {
                int pick = -1; double pa = -1;
                for (int k = 0; k < nk; ++k) {
                    const BNode &c = B.nodes[kids[k]];
                    if (c.obj < 0 && c.box.area() > pa) { pa = c.box.area(); pick = k; }
                }
                if (pick < 0) break;
                int c = kids[pick];
                kids[pick] = B.nodes[c].left;
                kids[nk++] = B.nodes[c].right;
            }
        }
        w.nk = nk;
        for (int k = 0; k < nk; ++k) {
            const BNode &c = B.nodes[kids[k]];
            for (int a = 0; a < 3; ++a) {
                double pad = pad_rel * (ext + std::fabs(c.box.lo[a]) + std::fabs(c.box.hi[a])) + 1e-30;
                w.lo[a][k] = down(c.box.lo[a], pad);
                w.hi[a][k] = up(c.box.hi[a], pad);
            }
            if (c.obj >= 0) w.ref[k] = leaf_ref[c.obj];
            else stack.push_back({kids[k], id, k, it.depth + 1});
        }
    };
    // the whole subtree of `root_item` into `W` (ids = positions in W, the root at W.size() on entry); returns the deepest level
    auto collapse = [&](const Item &root_item, std::vector<Wide> &W) {
        std::vector<Item> stack{root_item};
        int deepest = root_item.depth;
        bool first = true;
        while (!stack.empty()) {
            Item it = stack.back(); stack.pop_back();
            deepest = std::max(deepest, it.depth);
            const int id = (int)W.size();
            W.push_back(empty_node());
            if (!first) W[it.slot_node].ref[it.slot] = id;
            first = false;
            Wide w = W[id];
            expand(it, id, w, stack);
            W[id] = w;
        }
        return deepest;
    };
    std::vector<Wide> wide;
    out.root = 0;
    out.max_depth = 1;
    const int frontier = n >= (1 << 15) ? std::max(1 << 10, n >> 8) : 0;      // objects below a frontier subtree; 0: one thread
    if (!frontier) {
        wide.reserve((size_t)n / 2 + 1);
        out.max_depth = collapse({broot, -1, -1, 1}, wide);
    } else {
        struct Sub { Item it; std::vector<Wide> W; int deepest = 0; int base = 0; };
        struct Event { int top; int sub; };                    // pop order: a node of the top (index into `top`) or a frontier subtree
        std::vector<Wide> top;                                 // ids local to the top: refs between top nodes are patched below
        std::vector<Item> top_item;
        std::vector<Sub> subs;
        std::vector<Event> events;
        std::vector<Item> stack{{broot, -1, -1, 1}};
        while (!stack.empty()) {
            Item it = stack.back(); stack.pop_back();
            if (it.slot_node >= 0 && B.nodes[it.bnode].cnt <= frontier) {
                events.push_back({-1, (int)subs.size()});
                subs.emplace_back();
                subs.back().it = it;
                continue;
            }
            out.max_depth = std::max(out.max_depth, it.depth);
            const int id = (int)top.size();
            events.push_back({id, -1});
            top.push_back(empty_node());
            top_item.push_back(it);
            Wide w = top[id];
            expand(it, id, w, stack);
            top[id] = w;
        }
        parallel_for((int)subs.size(), host_threads(), [&](int i) {
            Sub &sb = subs[i];
            sb.W.reserve((size_t)B.nodes[sb.it.bnode].cnt / 2 + 1);
            Item root_item = sb.it;
            root_item.slot_node = -1;
            sb.deepest = collapse(root_item, sb.W);
        });
        std::vector<int> top_id(top.size());
        size_t total = 0;
        for (const Event &e : events) {
            if (e.top >= 0) top_id[e.top] = (int)total++;
            else { subs[e.sub].base = (int)total; total += subs[e.sub].W.size(); }
        }
        wide.assign(total, empty_node());
        for (size_t t = 0; t < top.size(); ++t) {
            wide[top_id[t]] = top[t];
            const Item &it = top_item[t];
            if (it.slot_node >= 0) wide[top_id[it.slot_node]].ref[it.slot] = top_id[t];      // (parents come first in pop order)
        }
        for (const Sub &sb : subs) {
            wide[top_id[sb.it.slot_node]].ref[sb.it.slot] = sb.base;
            out.max_depth = std::max(out.max_depth, sb.deepest);
        }
        parallel_for((int)subs.size(), host_threads(), [&](int i) {
            const Sub &sb = subs[i];
            for (size_t j = 0; j < sb.W.size(); ++j) {
                Wide w = sb.W[j];
                for (int k = 0; k < 4; ++k) if (w.ref[k] >= 0 && w.ref[k] != FAST_EMPTY) w.ref[k] += sb.base;
                wide[(size_t)sb.base + j] = w;
            }
        });
    }

    // Quantise: per node an origin p (float) and a power-of-two cell size 2^e per axis; child planes are bytes on that
    // grid.  The kernel turns a byte q into the float 32768 + q with one byte permute (rt_fast.cuh byte_f), so the
    // stored origin is p - 32768 cells and a plane is  stored_origin + (32768 + q) * cell.  In the kernel's ray-space
    // form of that sum the two large terms cancel, leaving roundings of up to ~cell/256; every plane is therefore
    // placed at least cell/64 outside the float box (lo down, hi up): the decoded box contains the float box with
    // that margin and culling stays conservative.
    out.nodes.assign(wide.size() * FAST_NODE_F4, make_float4(0, 0, 0, 0));
    std::atomic<bool> too_large(false);
    const int n_chunks = (int)((wide.size() + 4095) / 4096);       // nodes are independent: chunks of 4096 on the host's threads
    parallel_for(n_chunks, wide.size() >= (1u << 15) ? host_threads() : 1, [&](int chunk) {
    for (size_t id = (size_t)chunk * 4096; id < std::min(wide.size(), (size_t)(chunk + 1) * 4096); ++id) {
        const Wide &w = wide[id];
        float p[3]; uint32_t ebyte[3]; uint32_t qlo[3] = {0, 0, 0}, qhi[3] = {0, 0, 0};
        for (int a = 0; a < 3; ++a) {
            float mn = FLT_MAX, mx = -FLT_MAX;
            for (int k = 0; k < w.nk; ++k) { mn = std::min(mn, w.lo[a][k]); mx = std::max(mx, w.hi[a][k]); }
            if (w.nk == 0) { mn = mx = 0.f; }
            int e = -100;
            double extent = (double)mx - (double)mn;
            if (extent > 0) {
                e = (int)std::ceil(std::log2(extent / 254.0));
                if (e < -100) e = -100;
            }
            uint32_t lq[4], hq[4];
            for (;; ++e) {     // raise e until every plane, with its margin, fits in a byte
                if (e > 100) { too_large = true; return; }
                const double cell = std::ldexp(1.0, e), margin = cell / 64.0;
                const float pf = (float)((double)mn - cell / 32.0 - 32768.0 * cell);
                const double origin = (double)pf + 32768.0 * cell;      // where byte 0 decodes to
                bool ok = std::isfinite(pf);
                for (int k = 0; k < w.nk && ok; ++k) {
                    double l = std::floor(((double)w.lo[a][k] - margin - origin) / cell);
                    double h = std::ceil(((double)w.hi[a][k] + margin - origin) / cell);
                    while (origin + l * cell > (double)w.lo[a][k] - margin) l -= 1;
                    while (origin + h * cell < (double)w.hi[a][k] + margin) h += 1;
                    ok = l >= 0 && h <= 255;
                    lq[k] = (uint32_t)l; hq[k] = (uint32_t)h;
                }
                if (ok) { p[a] = pf; break; }
            }
            ebyte[a] = (uint32_t)(e + 127);
            for (int k = 0; k < 4; ++k) {
                uint32_t lo_q = 255, hi_q = 0;     // empty slot: inverted interval
                if (k < w.nk) { lo_q = lq[k]; hi_q = hq[k]; }
                qlo[a] |= lo_q << (8 * k); qhi[a] |= hi_q << (8 * k);
            }
        }
        float4 *nd = &out.nodes[id * FAST_NODE_F4];
        // cell sizes 2^e as float patterns ebyte << 23: x and y as upper halves packed in one word, z whole
        nd[0] = make_float4(p[0], p[1], p[2], i2f((int32_t)((ebyte[0] << 23) | (ebyte[1] << 7))));
        nd[1] = make_float4(i2f((int32_t)qlo[0]), i2f((int32_t)qlo[1]), i2f((int32_t)qlo[2]), i2f((int32_t)qhi[0]));
        nd[2] = make_float4(i2f((int32_t)qhi[1]), i2f((int32_t)qhi[2]), i2f((int32_t)FAST_NODE_BIAS_WORD), i2f((int32_t)(ebyte[2] << 23)));
        int32_t ref[4];
        for (int k = 0; k < 4; ++k) ref[k] = w.ref[k] == FAST_EMPTY ? ~(int32_t)(FAST_LEAF_SPHERE | n) : w.ref[k];
        nd[3] = make_float4(i2f(ref[0]), i2f(ref[1]), i2f(ref[2]), i2f(ref[3]));
    }
    });
    if (too_large) return fail("box extent too large to quantise");
    return true;
}

}  // namespace rtow
