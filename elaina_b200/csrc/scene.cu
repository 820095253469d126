// Scene ingest + SNCH-LBVH construction on the device, and the C ABI of the geometry queries.
//
// Build pipeline (replaces lbvh::bvh::construct, ext/lbvh/include/snch_lbvh/core/bvh.cuh:380-613, which is a
// chain of Thrust calls over four separate arrays):
//   leaf boxes + cones -> whole box (order-independent fmin/fmax reduction) -> 10-bit/axis Morton codes ->
//   stable LSD radix sort of (code, index) [cub::DeviceRadixSort] -> duplicate check -> 64-bit (code<<32|index)
//   keys when codes collide -> Karras topology -> bottom-up box refit -> bottom-up cone refit ->
//   fused traversal records (TNode/CNode) + Morton-ordered inline leaf records.
// Every floating-point step uses the reference's operation order (no FMA), so nodes, boxes and query results
// equal the host-compiled reference bit for bit; cones use detmath (see detmath.h).
// Silhouette adjacency (vertex/edge -> faces) is still assembled on the host like the reference
// (scene.cuh:631-676, 1135-1229), with a hash map instead of std::map; SURVEY 8(f) row N1 moves it to the device.
#include "scene.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cstring>
#include <algorithm>
#include <mutex>
#include <unordered_map>

namespace wost
{
    static thread_local std::string g_last_error;
    void set_last_error(const std::string &msg) { g_last_error = msg; }
    const std::string &last_error() { return g_last_error; }
    // ------------------------------------------------------------------ caching device allocator (common.cuh)
    namespace
    {
        struct AllocCache
        {
            std::mutex mu;
            std::unordered_multimap<unsigned long long, void *> free_blocks; // key = device << 48 | bytes
            size_t cached_bytes = 0;
        };
        AllocCache &alloc_cache()
        {
            static AllocCache c;
            return c;
        }
        // every block is cached (a scene build makes ~50 small allocations; cudaMalloc + cudaFree of each cost more than the kernels
        // that fill them): bounded by total bytes and by the number of parked blocks
        constexpr size_t kCacheMinBytes = 1, kCacheMaxBytes = (size_t)16 << 30, kCacheMaxBlocks = 8192;
        bool alloc_cache_enabled()
        {
            static const bool on = getenv("WOST_NO_ALLOC_CACHE") == nullptr;
            return on;
        }
        unsigned long long cache_key(int dev, size_t bytes) { return ((unsigned long long)dev << 48) | (unsigned long long)bytes; }
        int current_device()
        {
            int dev = 0;
            cudaGetDevice(&dev);
            return dev;
        }
    } // namespace
    cudaError_t dev_alloc(void **ptr, size_t bytes)
    {
        if (bytes >= kCacheMinBytes && alloc_cache_enabled())
        {
            AllocCache &c = alloc_cache();
            std::lock_guard<std::mutex> lock(c.mu);
            auto it = c.free_blocks.find(cache_key(current_device(), bytes));
            if (it != c.free_blocks.end())
            {
                *ptr = it->second;
                c.free_blocks.erase(it);
                c.cached_bytes -= bytes;
                return cudaSuccess;
            }
        }
        cudaError_t e = cudaMalloc(ptr, bytes);
        if (e != cudaSuccess && bytes >= kCacheMinBytes)
        {
            // out of memory with blocks parked in the cache: give them back and retry once
            cudaGetLastError();
            wost_trim_allocation_cache();
            e = cudaMalloc(ptr, bytes);
        }
        return e;
    }
    void dev_free(void *ptr, size_t bytes)
    {
        if (!ptr)
            return;
        if (bytes >= kCacheMinBytes && alloc_cache_enabled())
        {
            AllocCache &c = alloc_cache();
            std::lock_guard<std::mutex> lock(c.mu);
            if (c.cached_bytes + bytes <= kCacheMaxBytes && c.free_blocks.size() < kCacheMaxBlocks)
            {
                // The block is parked under the device that OWNS it, which need not be the current one (wost_set_device is
                // public: one process may drive several GPUs and destroy a handle while another device is current).
                cudaPointerAttributes attr;
                int owner = current_device();
                if (cudaPointerGetAttributes(&attr, ptr) == cudaSuccess && attr.type == cudaMemoryTypeDevice)
                    owner = attr.device;
                else
                    cudaGetLastError();
                // a parked block may be handed out again at once: no work of any stream of its device may still touch it
                // (cudaFree waits too)
                const int cur = current_device();
                if (owner != cur)
                    cudaSetDevice(owner);
                cudaDeviceSynchronize();
                if (owner != cur)
                    cudaSetDevice(cur);
                c.free_blocks.emplace(cache_key(owner, bytes), ptr);
                c.cached_bytes += bytes;
                return;
            }
        }
        cudaFree(ptr);
    }

    int sm_count()
    {
        int dev = 0, n = 148;
        if (cudaGetDevice(&dev) == cudaSuccess)
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        return n;
    }

    // ------------------------------------------------------------------ device helpers on raw arrays
    template <int D>
    WOST_DEV Vec<D> ld_vert(const float *verts, int i)
    {
        Vec<D> v;
#pragma unroll
        for (int k = 0; k < D; ++k)
            v[k] = verts[(size_t)i * D + k];
        return v;
    }
    template <int D>
    WOST_DEV Box<D> ld_box(const float *boxes, size_t i)
    {
        Box<D> b;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            b.up[k] = boxes[i * 2 * D + k];
            b.lo[k] = boxes[i * 2 * D + D + k];
        }
        return b;
    }
    template <int D>
    WOST_DEV Box<D> ld_box_cg(const float *boxes, size_t i) // L2-coherent read (written by another SM)
    {
        Box<D> b;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            b.up[k] = __ldcg(boxes + i * 2 * D + k);
            b.lo[k] = __ldcg(boxes + i * 2 * D + D + k);
        }
        return b;
    }
    template <int D>
    WOST_DEV void st_box(float *boxes, size_t i, const Box<D> &b)
    {
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            boxes[i * 2 * D + k] = b.up[k];
            boxes[i * 2 * D + D + k] = b.lo[k];
        }
    }
    template <int D>
    WOST_DEV Cone<D> ld_cone_cg(const float *cones, size_t i)
    {
        Cone<D> c;
#pragma unroll
        for (int k = 0; k < D; ++k)
            c.axis[k] = __ldcg(cones + i * (D + 2) + k);
        c.half_angle = __ldcg(cones + i * (D + 2) + D);
        c.radius = __ldcg(cones + i * (D + 2) + D + 1);
        return c;
    }
    template <int D>
    WOST_DEV void st_cone(float *cones, size_t i, const Cone<D> &c)
    {
#pragma unroll
        for (int k = 0; k < D; ++k)
            cones[i * (D + 2) + k] = c.axis[k];
        cones[i * (D + 2) + D] = c.half_angle;
        cones[i * (D + 2) + D + 1] = c.radius;
    }

    // silhouette record helpers (scene.cuh:322-353 / 743-787)
    template <int D>
    WOST_DEV bool sil_has_face(const int *rec, int f)
    {
        if (D == 2)
            return f == 0 ? rec[2] != -1 : rec[0] != -1;
        return f == 0 ? rec[3] != -1 : rec[0] != -1;
    }
    WOST_DEV Vec<2> sil_face_normal(const float *verts, const int *rec, int f, bool do_norm, Vec<2> *)
    {
        const int i = f == 0 ? 1 : 0;
        const Vec<2> d = vsub(ld_vert<2>(verts, rec[i + 1]), ld_vert<2>(verts, rec[i]));
        Vec<2> n;
        n[0] = d[1], n[1] = -d[0];
        return do_norm ? vnormalize(n) : n;
    }
    WOST_DEV Vec<3> sil_face_normal(const float *verts, const int *rec, int f, bool do_norm, Vec<3> *)
    {
        int i, j, k;
        if (f == 0)
            i = 3, j = 1, k = 2;
        else
            i = 0, j = 2, k = 1;
        const Vec<3> pa = ld_vert<3>(verts, rec[j]);
        const Vec<3> n = vcross(vsub(ld_vert<3>(verts, rec[k]), pa), vsub(ld_vert<3>(verts, rec[i]), pa));
        return do_norm ? vnormalize(n) : n;
    }
    template <int D>
    WOST_DEV Vec<D> sil_normal(const float *verts, const int *rec)
    {
        Vec<D> n;
#pragma unroll
        for (int k = 0; k < D; ++k)
            n[k] = 0.0f;
        for (int f = 0; f < 2; ++f)
            if (sil_has_face<D>(rec, f))
            {
                const Vec<D> nf = sil_face_normal(verts, rec, f, false, (Vec<D> *)nullptr);
#pragma unroll
                for (int k = 0; k < D; ++k)
                    n[k] = n[k] + nf[k];
            }
        return vnormalize(n);
    }
    WOST_DEV Vec<2> sil_centroid(const float *verts, const int *rec, Vec<2> *) { return ld_vert<2>(verts, rec[1]); }
    WOST_DEV Vec<3> sil_centroid(const float *verts, const int *rec, Vec<3> *)
    {
        const Vec<3> pa = ld_vert<3>(verts, rec[1]), pb = ld_vert<3>(verts, rec[2]);
        Vec<3> c;
        c[0] = dm_div(pa[0] + pb[0], 2.0f), c[1] = dm_div(pa[1] + pb[1], 2.0f), c[2] = dm_div(pa[2] + pb[2], 2.0f);
        return c;
    }

    // scene.cuh:420-433 / 870-885
    WOST_DEV Box<2> prim_box(const float *verts, const int *prim, Box<2> *)
    {
        const Vec<2> p0 = ld_vert<2>(verts, prim[0]), p1 = ld_vert<2>(verts, prim[1]);
        Box<2> b;
#pragma unroll
        for (int k = 0; k < 2; ++k)
        {
            b.up[k] = dm_max(p0[k], p1[k]) + kBvhOffset;
            b.lo[k] = dm_min(p0[k], p1[k]) - kBvhOffset;
        }
        return b;
    }
    WOST_DEV Box<3> prim_box(const float *verts, const int *prim, Box<3> *)
    {
        const Vec<3> pa = ld_vert<3>(verts, prim[0]);
        Box<3> b;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            b.up[k] = pa[k] + kEps;
            b.lo[k] = pa[k] - kEps;
        }
        for (int v = 1; v < 3; ++v)
        {
            const Vec<3> p = ld_vert<3>(verts, prim[v]);
#pragma unroll
            for (int k = 0; k < 3; ++k)
            {
                b.lo[k] = dm_fmin_strict(b.lo[k], p[k] - kEps);
                b.up[k] = dm_fmax_strict(b.up[k], p[k] + kEps);
            }
        }
        return b;
    }

    // exact (unpadded) extents: the reference's padded box is fl(extent +- pad) per coordinate, for leaves AND for
    // merged internal boxes (max_i fl(a_i + c) == fl(max_i a_i + c)), so the traversal records store the exact extents
    // and rebuild the padded box with one add per coordinate (lbvh.cuh: pad_box).
    template <int D>
    WOST_DEV Box<D> prim_extent(const float *verts, const int *prim)
    {
        Box<D> b;
        const Vec<D> p0 = ld_vert<D>(verts, prim[0]);
#pragma unroll
        for (int k = 0; k < D; ++k)
            b.up[k] = b.lo[k] = p0[k];
        for (int v = 1; v < D; ++v)
        {
            const Vec<D> p = ld_vert<D>(verts, prim[v]);
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                b.lo[k] = dm_fmin_strict(b.lo[k], p[k]);
                b.up[k] = dm_fmax_strict(b.up[k], p[k]);
            }
        }
        return b;
    }

    // ------------------------------------------------------------------ build kernels
    // leaf boxes + cones in ORIGINAL primitive order (scene.cuh:435-502 / 887-961)
    template <int D>
    __global__ void k_leaf_boxes_cones(int n, const float *verts, const int *prims, const int *sil_rec, const int *prim_sil,
                                       float *lboxes, float *lcones)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const Box<D> box = prim_box(verts, prims + (size_t)i * D, (Box<D> *)nullptr);
        st_box<D>(lboxes, i, box);
        const Vec<D> bc = box_centroid(box);
        Cone<D> c;
#pragma unroll
        for (int k = 0; k < D; ++k)
            c.axis[k] = 0.0f;
        c.half_angle = kPi;
        c.radius = 0.0f;
        bool any = false, all_two = true;
        for (int j = 0; j < D; ++j)
        {
            const int si = prim_sil[(size_t)i * D + j];
            if (si == -1)
                continue;
            const int *rec = sil_rec + (size_t)si * 4;
            const Vec<D> nn = sil_normal<D>(verts, rec);
#pragma unroll
            for (int k = 0; k < D; ++k)
                c.axis[k] += nn[k];
            c.radius = dm_max(c.radius, vlen(vsub(sil_centroid(verts, rec, (Vec<D> *)nullptr), bc)));
            all_two = all_two && sil_has_face<D>(rec, 0) && sil_has_face<D>(rec, 1);
            any = true;
        }
        if (!any)
            c.half_angle = -kPi;
        else if (!all_two)
            c.half_angle = kPi;
        else
        {
            const float an = vlen(c.axis);
            if (an > kEps)
            {
#pragma unroll
                for (int k = 0; k < D; ++k)
                    c.axis[k] = dm_div(c.axis[k], an);
                c.half_angle = 0.0f;
                for (int j = 0; j < D; ++j)
                {
                    const int si = prim_sil[(size_t)i * D + j];
                    if (si == -1)
                        continue;
                    const int *rec = sil_rec + (size_t)si * 4;
                    for (int f = 0; f < 2; ++f)
                    {
                        Vec<D> nn;
                        if (sil_has_face<D>(rec, f))
                            nn = sil_face_normal(verts, rec, f, true, (Vec<D> *)nullptr);
                        else
                        {
#pragma unroll
                            for (int k = 0; k < D; ++k)
                                nn[k] = 0.0f;
                        }
                        const float ang = dm_acos(dm_max(-1.0f, dm_min(1.0f, vdot(c.axis, nn))));
                        c.half_angle = dm_max(c.half_angle, ang);
                    }
                }
            }
        }
        st_cone<D>(lcones, i, c);
    }

    // whole box: per-block partials, then one block (fmin/fmax are associative+commutative: result is order-independent)
    template <int D>
    __global__ void k_reduce_boxes(int n, const float *boxes, float *partial)
    {
        __shared__ float sh[2 * D][256];
        float up[D], lo[D];
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            up[k] = -dm_inf();
            lo[k] = dm_inf();
        }
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        {
            const Box<D> b = ld_box<D>(boxes, i);
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                up[k] = dm_fmax_strict(up[k], b.up[k]);
                lo[k] = dm_fmin_strict(lo[k], b.lo[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            sh[k][threadIdx.x] = up[k];
            sh[D + k][threadIdx.x] = lo[k];
        }
        __syncthreads();
        for (int s = 128; s > 0; s >>= 1)
        {
            if (threadIdx.x < s)
            {
#pragma unroll
                for (int k = 0; k < D; ++k)
                {
                    sh[k][threadIdx.x] = dm_fmax_strict(sh[k][threadIdx.x], sh[k][threadIdx.x + s]);
                    sh[D + k][threadIdx.x] = dm_fmin_strict(sh[D + k][threadIdx.x], sh[D + k][threadIdx.x + s]);
                }
            }
            __syncthreads();
        }
        if (threadIdx.x == 0)
        {
#pragma unroll
            for (int k = 0; k < 2 * D; ++k)
                partial[(size_t)blockIdx.x * 2 * D + k] = sh[k][0];
        }
    }

    WOST_DEV uint32_t expand_bits(uint32_t v) // core/morton_code.cuh:19-38
    {
        v = (v | (v << 16)) & 0x070000FFu;
        v = (v | (v << 8)) & 0x0700F00Fu;
        v = (v | (v << 4)) & 0x430C30C3u;
        v = (v | (v << 2)) & 0x49249249u;
        return v;
    }
    // core/bvh.cuh:258-304 + core/morton_code.cuh:61-70,85-92
    template <int D>
    __global__ void k_morton(int n, const float *lboxes, const float *whole, uint32_t *code, uint32_t *index)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const Box<D> b = ld_box<D>(lboxes, i);
        const Vec<D> p = box_centroid(b);
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            const float wl = whole[D + k], wu = whole[k];
            float x = p[k] - wl;
            x = dm_div(x, wu - wl);
            x = dm_fmin_strict(dm_fmax_strict(x * 1024.0f, 0.0f), 1024.0f - 1.0f);
            const uint32_t e = expand_bits((uint32_t)x);
            if (D == 2)
                c = (k == 0) ? e * 2 : c + e;
            else
                c = (k == 0) ? e * 4 : ((k == 1) ? c + e * 2 : c + e);
        }
        code[i] = c;
        index[i] = (uint32_t)i;
    }
    __global__ void k_check_unique(int n, const uint32_t *code, int *dup)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n - 1)
            return;
        if (code[i] == code[i + 1])
            *dup = 1;
    }
    __global__ void k_keys64(int n, const uint32_t *code, const uint32_t *order, unsigned long long *key)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n)
            key[i] = ((unsigned long long)code[i] << 32) | (unsigned long long)order[i];
    }
    WOST_DEV int cub_bits(uint32_t a, uint32_t b) { return __clz(a ^ b); }
    WOST_DEV int cub_bits(unsigned long long a, unsigned long long b) { return __clzll(a ^ b); }
    // core/bvh.cuh:110-229 (Karras 2012)
    template <typename U>
    __global__ void k_karras(int n, const U *code, Node *nodes, int *leaf_count, int *leaf_first)
    {
        const int idx = blockIdx.x * blockDim.x + threadIdx.x;
        if (idx >= n - 1)
            return;
        int first, last;
        if (idx == 0)
        {
            first = 0;
            last = n - 1;
        }
        else
        {
            const U self = code[idx];
            const int Ld = cub_bits(self, code[idx - 1]);
            const int Rd = cub_bits(self, code[idx + 1]);
            const int d = (Rd > Ld) ? 1 : -1;
            const int dmin = min(Ld, Rd);
            int lmax = 2;
            int delta = -1;
            int it = idx + d * lmax;
            if (0 <= it && it < n)
                delta = cub_bits(self, code[it]);
            while (delta > dmin)
            {
                lmax <<= 1;
                it = idx + d * lmax;
                delta = -1;
                if (0 <= it && it < n)
                    delta = cub_bits(self, code[it]);
            }
            int l = 0;
            int t = lmax >> 1;
            while (t > 0)
            {
                it = idx + (l + t) * d;
                delta = -1;
                if (0 <= it && it < n)
                    delta = cub_bits(self, code[it]);
                if (delta > dmin)
                    l += t;
                t >>= 1;
            }
            const int j = idx + l * d;
            first = min(idx, j);
            last = max(idx, j);
        }
        int gamma;
        {
            const U fc = code[first], lc = code[last];
            if (fc == lc)
                gamma = (first + last) >> 1;
            else
            {
                const int dn = cub_bits(fc, lc);
                int split = first;
                int stride = last - first;
                do
                {
                    stride = (stride + 1) >> 1;
                    const int mid = split + stride;
                    if (mid < last)
                    {
                        const int delta = cub_bits(fc, code[mid]);
                        if (delta > dn)
                            split = mid;
                    }
                } while (stride > 1);
                gamma = split;
            }
        }
        uint32_t L = gamma, R = gamma + 1;
        if (first == gamma)
            L += n - 1;
        if (last == gamma + 1)
            R += n - 1;
        leaf_count[idx] = last - first + 1;
        leaf_first[idx] = first;
        nodes[idx].left = L;
        nodes[idx].right = R;
        nodes[idx].object = kInvalid;
        nodes[L].parent = idx;
        nodes[R].parent = idx;
        if (idx == 0)
            nodes[0].parent = kInvalid;
    }
    // leaves in sorted order: object index, box, cone (core/bvh.cuh:452-499)
    template <int D>
    __global__ void k_init_leaves(int n, const uint32_t *order, const float *lboxes, const float *lcones, Node *nodes, float *boxes, float *cones)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const uint32_t o = order[i];
        const size_t dst = (size_t)(n - 1) + i;
        nodes[dst].left = kInvalid;
        nodes[dst].right = kInvalid;
        nodes[dst].object = o;
        if (n == 1)
            nodes[dst].parent = kInvalid;
#pragma unroll
        for (int k = 0; k < 2 * D; ++k)
            boxes[dst * 2 * D + k] = lboxes[(size_t)o * 2 * D + k];
#pragma unroll
        for (int k = 0; k < D + 2; ++k)
            cones[dst * (D + 2) + k] = lcones[(size_t)o * (D + 2) + k];
    }
    // bottom-up refits (core/bvh.cuh:518-604): the second thread to reach a node merges its children.
    template <int D>
    __global__ void k_refit_boxes(int n, const Node *nodes, float *boxes, int *flags)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        uint32_t parent = nodes[(size_t)(n - 1) + i].parent;
        while (parent != kInvalid)
        {
            __threadfence();
            const int old = atomicAdd(flags + parent, 1);
            if (old == 0)
                return;
            const Box<D> lb = ld_box_cg<D>(boxes, nodes[parent].left);
            const Box<D> rb = ld_box_cg<D>(boxes, nodes[parent].right);
            st_box<D>(boxes, parent, box_merge(lb, rb));
            parent = nodes[parent].parent;
        }
    }
    template <int D>
    __global__ void k_refit_cones(int n, const Node *nodes, const float *boxes, float *cones, int *flags)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        uint32_t parent = nodes[(size_t)(n - 1) + i].parent;
        while (parent != kInvalid)
        {
            __threadfence();
            const int old = atomicAdd(flags + parent, 1);
            if (old == 0)
                return;
            const uint32_t l = nodes[parent].left, r = nodes[parent].right;
            const Cone<D> lc = ld_cone_cg<D>(cones, l), rc = ld_cone_cg<D>(cones, r);
            const Box<D> lb = ld_box<D>(boxes, l), rb = ld_box<D>(boxes, r), pb = ld_box<D>(boxes, parent);
            st_cone<D>(cones, parent, cone_merge(lc, rc, box_centroid(lb), box_centroid(rb), box_centroid(pb)));
            parent = nodes[parent].parent;
        }
    }

    // exact extents of the leaves in sorted order (internal nodes: k_refit_boxes on this buffer)
    template <int D>
    __global__ void k_leaf_extents(int n, const uint32_t *order, const float *verts, const int *prims, float *ext)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        st_box<D>(ext, (size_t)(n - 1) + i, prim_extent<D>(verts, prims + (size_t)order[i] * D));
    }
    // Visit rank of every leaf in the reference's closest-point traversal order (core/query.cuh:279-314): at a node the
    // L leaf child is evaluated first, then the R leaf child, then the whole R subtree (pushed last, popped first), then
    // the L subtree. Among exactly tied distances the reference keeps the LAST visited leaf ("<="), i.e. the one with the
    // highest rank -- which lets query_nearest visit nodes in any order and still return the reference's index.
    __global__ void k_leaf_rank(int n, const Node *nodes, const int *leaf_count, uint32_t *rank)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const uint32_t first_leaf = (uint32_t)(n - 1);
        uint32_t child = first_leaf + i, r = 0;
        uint32_t p = nodes[child].parent;
        while (p != kInvalid)
        {
            const uint32_t L = nodes[p].left, R = nodes[p].right;
            const bool Lleaf = L >= first_leaf, Rleaf = R >= first_leaf;
            if (child == L)
            {
                if (!Lleaf)
                    r += Rleaf ? 1u : (uint32_t)leaf_count[R];
            }
            else
                r += Lleaf ? 1u : 0u;
            child = p;
            p = nodes[p].parent;
        }
        rank[i] = r;
    }


    // ------------------------------------------------------------------ oriented boxes of the traversal records (lbvh.cuh: Obb)
    constexpr int kObbFloats2 = 6, kObbFloats3 = 12;
    template <int D>
    constexpr int obb_floats() { return D == 2 ? kObbFloats2 : kObbFloats3; }
    WOST_DEV Vec<2> prim_weighted_normal(const LeafPrim<2> &p)
    {
        Vec<2> n;
        n[0] = p.p[1] - p.p[3], n[1] = p.p[2] - p.p[0]; // perpendicular of pb - pa, length = segment length
        return n;
    }
    WOST_DEV Vec<3> prim_weighted_normal(const LeafPrim<3> &p)
    {
        Vec<3> a, b;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            a[k] = p.p[3 + k] - p.p[k], b[k] = p.p[6 + k] - p.p[k];
        return vcross(a, b); // length = 2 x area
    }
    // (1) frame: length / area weighted normals summed bottom-up (the second thread to reach a node merges its children,
    // like the refits). The right child's sum is flipped when it opposes the left one: stroke soups are not consistently
    // oriented, and the sign of a slab normal does not matter.
    template <int D>
    __global__ void k_obb_normals(int n, const Node *nodes, const LeafPrim<D> *lp, float *nsum, int *flags)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const Vec<D> w = prim_weighted_normal(lp[i]);
        const size_t leaf = (size_t)(n - 1) + i;
#pragma unroll
        for (int k = 0; k < D; ++k)
            nsum[leaf * D + k] = w[k];
        uint32_t parent = nodes[leaf].parent;
        while (parent != kInvalid)
        {
            __threadfence();
            const int old = atomicAdd(flags + parent, 1);
            if (old == 0)
                return;
            const size_t l = nodes[parent].left, r = nodes[parent].right;
            float a[D], b[D], dot = 0.0f;
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                a[k] = __ldcg(nsum + l * D + k);
                b[k] = __ldcg(nsum + r * D + k);
                dot = dot + a[k] * b[k];
            }
#pragma unroll
            for (int k = 0; k < D; ++k)
                nsum[(size_t)parent * D + k] = (dot >= 0.0f) ? a[k] + b[k] : a[k] - b[k];
            parent = nodes[parent].parent;
        }
    }
    WOST_DEV Obb<2> obb_frame(const float *ns, const LeafPrim<2> *)
    {
        Obb<2> o;
        const float len = dm_sqrt(ns[0] * ns[0] + ns[1] * ns[1]);
        if (len > 0.0f && len < dm_inf())
            o.ax[0] = dm_div(ns[0], len), o.ax[1] = dm_div(ns[1], len);
        else
            o.ax[0] = 1.0f, o.ax[1] = 0.0f;
        return o;
    }
    // 3D: n = normalised sum; in-plane axis t = the longest edge (single triangle) or the coordinate axis least aligned with n,
    // made orthogonal to n (Gram-Schmidt)
    WOST_DEV Obb<3> obb_frame(const float *ns, const LeafPrim<3> *single)
    {
        Obb<3> o;
        const float len = dm_sqrt(ns[0] * ns[0] + ns[1] * ns[1] + ns[2] * ns[2]);
        if (len > 0.0f && len < dm_inf())
            o.n[0] = dm_div(ns[0], len), o.n[1] = dm_div(ns[1], len), o.n[2] = dm_div(ns[2], len);
        else
            o.n[0] = 0.0f, o.n[1] = 0.0f, o.n[2] = 1.0f;
        float e[3] = {0.0f, 0.0f, 0.0f};
        bool have = false;
        if (single)
        {
            float bestl = 0.0f;
            for (int a = 0; a < 3; ++a)
            {
                const int b = (a + 1) % 3;
                const float dx = single->p[3 * b] - single->p[3 * a], dy = single->p[3 * b + 1] - single->p[3 * a + 1], dz = single->p[3 * b + 2] - single->p[3 * a + 2];
                const float l2 = dx * dx + dy * dy + dz * dz;
                if (l2 > bestl && l2 < dm_inf())
                    bestl = l2, e[0] = dx, e[1] = dy, e[2] = dz, have = true;
            }
        }
        for (int attempt = 0; attempt < 2; ++attempt)
        {
            if (!have)
            {
                const float ax = dm_abs(o.n[0]), ay = dm_abs(o.n[1]), az = dm_abs(o.n[2]);
                const int k = (ax <= ay && ax <= az) ? 0 : (ay <= az ? 1 : 2);
                e[0] = e[1] = e[2] = 0.0f;
                e[k] = 1.0f;
            }
            const float dn = e[0] * o.n[0] + e[1] * o.n[1] + e[2] * o.n[2];
            const float tx = e[0] - dn * o.n[0], ty = e[1] - dn * o.n[1], tz = e[2] - dn * o.n[2];
            const float tl = dm_sqrt(tx * tx + ty * ty + tz * tz);
            if (tl > 1e-20f && tl < dm_inf() && (have ? tl * tl > 0.25f * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]) : true))
            {
                o.t[0] = dm_div(tx, tl), o.t[1] = dm_div(ty, tl), o.t[2] = dm_div(tz, tl);
                return o;
            }
            have = false; // degenerate triangle: fall back to the coordinate axis
        }
        o.t[0] = 1.0f, o.t[1] = 0.0f, o.t[2] = 0.0f; // unreachable for a unit n
        return o;
    }
    // (2) extents: G threads (1, a warp, a CTA) per node project every vertex of the node's leaf range onto the frame
    // (coalesced reads of the Morton-ordered leaf records) and keep the exact fp32 min / max, widened by 2^-20 max|coordinate|.
    // Nodes are dealt to the three variants by the size of their range (lo_count < count <= hi_count).
    template <int D, int G>
    __global__ void __launch_bounds__(256) k_obb_extents(int n, const int *leaf_first, const int *leaf_count, const LeafPrim<D> *lp, const float *nsum,
                                                                      const float *ext, float *obb, int lo_count, int hi_count)
    {
        constexpr int F = D == 2 ? kObbFloats2 : kObbFloats3;
        const int NN = 2 * n - 1;
        long long node;
        int lane;
        if (G == 1)
            node = (long long)blockIdx.x * blockDim.x + threadIdx.x, lane = 0;
        else if (G == 32)
            node = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
        else
            node = blockIdx.x, lane = threadIdx.x;
        if (node >= (G == 1 ? NN : n - 1))
            return;
        const bool is_leaf = node >= n - 1;
        const int first = is_leaf ? (int)(node - (n - 1)) : leaf_first[node];
        const int count = is_leaf ? 1 : leaf_count[node];
        if (count <= lo_count || count > hi_count)
            return;
        float ns[D];
#pragma unroll
        for (int k = 0; k < D; ++k)
            ns[k] = nsum[(size_t)node * D + k];
        Obb<D> o = obb_frame(ns, is_leaf ? lp + first : (const LeafPrim<D> *)nullptr);
        float mn[D], mx[D];
#pragma unroll
        for (int k = 0; k < D; ++k)
            mn[k] = dm_inf(), mx[k] = -dm_inf();
        for (int j = lane; j < count; j += G)
        {
            const LeafPrim<D> p = lp[(size_t)first + j];
#pragma unroll
            for (int v = 0; v < D; ++v)
            {
                Vec<D> x;
#pragma unroll
                for (int k = 0; k < D; ++k)
                    x[k] = p.p[v * D + k];
                float pr[D];
                obb_project(o, x, pr);
#pragma unroll
                for (int k = 0; k < D; ++k)
                    mn[k] = fmin_nn(mn[k], pr[k]), mx[k] = fmax_nn(mx[k], pr[k]);
            }
        }
        if constexpr (G > 1)
        {
#pragma unroll
            for (int k = 0; k < D; ++k)
#pragma unroll
                for (int off = 16; off > 0; off >>= 1)
                {
                    mn[k] = fmin_nn(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], off));
                    mx[k] = fmax_nn(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], off));
                }
        }
        if constexpr (G > 32)
        {
            __shared__ float sm[2 * D][G / 32];
            if ((threadIdx.x & 31) == 0)
#pragma unroll
                for (int k = 0; k < D; ++k)
                    sm[k][threadIdx.x >> 5] = mn[k], sm[D + k][threadIdx.x >> 5] = mx[k];
            __syncthreads();
            if (threadIdx.x == 0)
                for (int w = 1; w < G / 32; ++w)
#pragma unroll
                    for (int k = 0; k < D; ++k)
                        mn[k] = fmin_nn(mn[k], sm[k][w]), mx[k] = fmax_nn(mx[k], sm[D + k][w]);
        }
        if (lane != 0)
            return;
        float m = 0.0f;
#pragma unroll
        for (int k = 0; k < 2 * D; ++k)
            m = fmax_nn(m, dm_abs(ext[k])); // node 0 = exact extents of the whole mesh
        const float wdn = m * 9.5367431640625e-07f; // 2^-20
        float *dst = obb + (size_t)node * F;
        if constexpr (D == 2)
        {
            const Obb<2> &q = o;
            dst[0] = q.ax[0], dst[1] = q.ax[1];
            dst[2] = mn[0] - wdn, dst[3] = mx[0] + wdn, dst[4] = mn[1] - wdn, dst[5] = mx[1] + wdn;
        }
        else
        {
            const Obb<3> &q = o;
#pragma unroll
            for (int k = 0; k < 3; ++k)
                dst[k] = q.n[k], dst[3 + k] = q.t[k], dst[6 + k] = mn[k] - wdn, dst[9 + k] = mx[k] + wdn;
        }
    }

    // ------------------------------------------------------------------ record order: top levels first
    // The fused records are stored in breadth-first order for the top levels of the tree (depth < top_depth) and in Karras
    // order below: the records every query reads -- the first ~17 of the ~20 levels of a descent -- form one contiguous slab
    // at the start of the array, which the integrators pin in L2 with an access-policy window when the whole tree does not
    // fit (integrator.cu: apply_l2_window; c4: 280 MB of records against 126 MB of L2). Canonical arrays keep Karras order.
    __global__ void k_node_depth(int n, const Node *nodes, int *depth, int *hist)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n - 1)
            return;
        int d = 0;
        for (uint32_t p = nodes[i].parent; p != kInvalid; p = nodes[p].parent)
            ++d;
        depth[i] = d;
        atomicAdd(hist + min(d, 63), 1);
    }
    __global__ void k_bfs_keys(int n, const int *depth, int top_depth, unsigned long long *key, uint32_t *val)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n - 1)
            return;
        key[i] = ((unsigned long long)min(depth[i], top_depth) << 32) | (unsigned long long)(uint32_t)i;
        val[i] = (uint32_t)i;
    }
    __global__ void k_perm_scatter(int n, const uint32_t *sorted_old, uint32_t *perm)
    {
        const int j = blockIdx.x * blockDim.x + threadIdx.x;
        if (j >= n - 1)
            return;
        perm[sorted_old[j]] = (uint32_t)j;
    }

    // ------------------------------------------------------------------ traversal layout
    WOST_DEV uint32_t child_ref(const Node *nodes, uint32_t child, int n)
    {
        return (nodes[child].object != kInvalid) ? ((child - (uint32_t)(n - 1)) | kLeafFlag) : child;
    }
    template <int D>
    __global__ void k_layout_internal(int n, const Node *nodes, const float *ext, const float *cones, const float *obb, const uint32_t *perm, TNode<D> *tn, CNode<D> *cn)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n - 1)
            return;
        const uint32_t l = nodes[i].left, r = nodes[i].right;
        TNode<D> t;
#pragma unroll
        for (int k = 0; k < 2 * D; ++k)
        {
            t.lb[k] = ext[(size_t)l * 2 * D + k];
            t.rb[k] = ext[(size_t)r * 2 * D + k];
        }
        t.lref = child_ref(nodes, l, n);
        t.rref = child_ref(nodes, r, n);
        if (!(t.lref & kLeafFlag))
            t.lref = perm[t.lref];
        if (!(t.rref & kLeafFlag))
            t.rref = perm[t.rref];
        const uint32_t dst = perm[i];
        {
            constexpr int F = D == 2 ? kObbFloats2 : kObbFloats3;
            float *raw = reinterpret_cast<float *>(&t);
            constexpr int l0 = D == 2 ? 10 : 16; // float offset of the left child's oriented box inside the record
#pragma unroll
            for (int k = 0; k < F; ++k)
            {
                raw[l0 + k] = obb[(size_t)l * F + k];
                raw[l0 + F + k] = obb[(size_t)r * F + k];
            }
            if (D == 2)
                raw[22] = raw[23] = 0.0f;
            else
                raw[14] = raw[15] = 0.0f;
        }
        tn[dst] = t;
        CNode<D> c;
#pragma unroll
        for (int k = 0; k < D + 2; ++k)
        {
            c.lc[k] = cones[(size_t)l * (D + 2) + k];
            c.rc[k] = cones[(size_t)r * (D + 2) + k];
        }
        if (D == 3)
        {
            float *pad = reinterpret_cast<float *>(&c) + 10;
            pad[0] = 0.0f;
            pad[1] = 0.0f;
        }
        cn[dst] = c;
    }
    WOST_DEV void fill_sil(LeafSil<2> &o, const float *verts, const int *rec)
    {
        const Vec<2> p = ld_vert<2>(verts, rec[1]);
        o.p[0] = p[0], o.p[1] = p[1];
        const bool f0 = sil_has_face<2>(rec, 0), f1 = sil_has_face<2>(rec, 1);
        Vec<2> n0, n1;
        n0[0] = n0[1] = n1[0] = n1[1] = 0.0f;
        if (f0 && f1)
        {
            n0 = sil_face_normal(verts, rec, 0, true, (Vec<2> *)nullptr);
            n1 = sil_face_normal(verts, rec, 1, true, (Vec<2> *)nullptr);
        }
        o.n0[0] = n0[0], o.n0[1] = n0[1], o.n1[0] = n1[0], o.n1[1] = n1[1];
        o.flags = 1u | (f0 ? 2u : 0u) | (f1 ? 4u : 0u);
        o.pad = 0;
    }
    WOST_DEV void fill_sil(LeafSil<3> &o, const float *verts, const int *rec)
    {
        const Vec<3> pa = ld_vert<3>(verts, rec[1]), pb = ld_vert<3>(verts, rec[2]);
        const bool f0 = sil_has_face<3>(rec, 0), f1 = sil_has_face<3>(rec, 1);
        Vec<3> n0, n1;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            n0[k] = n1[k] = 0.0f;
        if (f0 && f1)
        {
            n0 = sil_face_normal(verts, rec, 0, true, (Vec<3> *)nullptr);
            n1 = sil_face_normal(verts, rec, 1, true, (Vec<3> *)nullptr);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            o.pa[k] = pa[k];
            o.pb[k] = pb[k];
            o.n0[k] = n0[k];
            o.n1[k] = n1[k];
        }
        o.flags = 1u | (f0 ? 2u : 0u) | (f1 ? 4u : 0u);
        o.pad[0] = o.pad[1] = o.pad[2] = 0;
    }
    WOST_DEV void fill_prim(LeafPrim<2> &o, const float *verts, const int *prim, uint32_t)
    {
        o.p[0] = verts[(size_t)prim[0] * 2], o.p[1] = verts[(size_t)prim[0] * 2 + 1];
        o.p[2] = verts[(size_t)prim[1] * 2], o.p[3] = verts[(size_t)prim[1] * 2 + 1];
    }
    WOST_DEV void fill_prim(LeafPrim<3> &o, const float *verts, const int *prim, uint32_t obj)
    {
#pragma unroll
        for (int v = 0; v < 3; ++v)
#pragma unroll
            for (int k = 0; k < 3; ++k)
                o.p[v * 3 + k] = verts[(size_t)prim[v] * 3 + k];
        o.obj = obj;
        o.pad[0] = o.pad[1] = 0.0f;
    }
    template <int D>
    __global__ void k_layout_leaves(int n, const uint32_t *order, const float *verts, const int *prims, const int *sil_rec,
                                    const int *prim_sil, LeafPrim<D> *lp, uint32_t *lobj, LeafSil<D> *ls)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        const uint32_t o = order[i];
        lobj[i] = o;
        LeafPrim<D> p;
        fill_prim(p, verts, prims + (size_t)o * D, o);
        lp[i] = p;
        for (int j = 0; j < D; ++j)
        {
            LeafSil<D> s;
            const int si = prim_sil[(size_t)o * D + j];
            if (si == -1)
            {
                float *raw = reinterpret_cast<float *>(&s);
                for (int k = 0; k < (int)(sizeof(LeafSil<D>) / 4); ++k)
                    raw[k] = 0.0f;
            }
            else
                fill_sil(s, verts, sil_rec + (size_t)si * 4);
            ls[(size_t)i * D + j] = s;
        }
    }

    // ------------------------------------------------------------------ host: silhouette adjacency
    static void host_silhouettes_2d(int nverts, int nprims, const int *prims, std::vector<int> &sil_rec, std::vector<int> &prim_sil)
    {
        sil_rec.assign((size_t)nverts * 4, -1); // scene.cuh:631-652
        for (int i = 0; i < nprims; ++i)
        {
            const int x = prims[2 * i], y = prims[2 * i + 1];
            sil_rec[(size_t)x * 4 + 1] = x;
            sil_rec[(size_t)x * 4 + 2] = y;
            sil_rec[(size_t)y * 4 + 0] = x;
            sil_rec[(size_t)y * 4 + 1] = y;
        }
        std::vector<char> seen(nverts, 0); // scene.cuh:654-672
        prim_sil.assign((size_t)nprims * 2, -1);
        for (int i = 0; i < nprims; ++i)
        {
            int ptr = 0;
            for (int j = 0; j < 2; ++j)
            {
                const int v = prims[2 * i + j];
                if (!seen[v])
                {
                    seen[v] = 1;
                    prim_sil[(size_t)i * 2 + ptr++] = v;
                }
            }
        }
    }
    static void host_silhouettes_3d(int nprims, const int *prims, std::vector<int> &sil_rec, std::vector<int> &prim_sil)
    {
        // scene.cuh:1135-1166: edge ids in first-seen order (hash map instead of std::map: same ids, O(1) lookups)
        std::unordered_map<unsigned long long, int> index_map;
        index_map.reserve((size_t)nprims * 2);
        std::vector<int> eot((size_t)nprims * 3);
        int E = 0;
        for (int i = 0; i < nprims; ++i)
            for (int j = 0; j < 3; ++j)
            {
                int I = prims[3 * i + j], J = prims[3 * i + (j + 1) % 3];
                if (I > J)
                    std::swap(I, J);
                const unsigned long long key = ((unsigned long long)(uint32_t)I << 32) | (uint32_t)J;
                auto it = index_map.find(key);
                if (it == index_map.end())
                    it = index_map.emplace(key, E++).first;
                eot[(size_t)i * 3 + j] = it->second;
            }
        sil_rec.assign((size_t)E * 4, -1); // scene.cuh:1167-1204
        for (int i = 0; i < nprims; ++i)
            for (int j = 0; j < 3; ++j)
            {
                const int I = j - 1 < 0 ? 2 : j - 1;
                int J = j, K = j + 1 > 2 ? 0 : j + 1;
                int orientation = 1;
                if (prims[3 * i + J] > prims[3 * i + K])
                {
                    std::swap(J, K);
                    orientation = -1;
                }
                int *rec = &sil_rec[(size_t)eot[(size_t)i * 3 + j] * 4];
                rec[orientation == 1 ? 0 : 3] = prims[3 * i + I];
                rec[1] = prims[3 * i + J];
                rec[2] = prims[3 * i + K];
            }
        std::vector<char> seen(E, 0); // scene.cuh:1205-1225
        prim_sil.assign((size_t)nprims * 3, -1);
        for (int i = 0; i < nprims; ++i)
        {
            int ptr = 0;
            for (int j = 0; j < 3; ++j)
            {
                const int e = eot[(size_t)i * 3 + j];
                if (!seen[e])
                {
                    seen[e] = 1;
                    prim_sil[(size_t)i * 3 + ptr++] = e;
                }
            }
        }
    }

    // ------------------------------------------------------------------ device: edge adjacency of a triangle mesh
    // Replaces the reference's host loops over std::map / std::unordered_map (scene.cuh:1135-1229; seconds at 10^6
    // triangles) with sort-by-edge-key. Slot = 3 * triangle + local edge. Same results as the sequential code:
    //   edge ids     in first-seen order            -> rank of the edge's first slot among all first slots (scan)
    //   record write "last triangle wins" per field -> the slots of one edge are replayed in ascending order by one thread
    //   ownership    first triangle that references -> the edge's first slot
    __global__ void k_edge_keys(int nprims, const int *prims, unsigned long long *keys, int *slots)
    {
        const int sl = blockIdx.x * blockDim.x + threadIdx.x;
        if (sl >= 3 * nprims)
            return;
        const int i = sl / 3, j = sl - 3 * i;
        int I = prims[3 * i + j], J = prims[3 * i + (j + 1) % 3];
        if (I > J)
        {
            const int t = I;
            I = J;
            J = t;
        }
        keys[sl] = ((unsigned long long)(uint32_t)I << 32) | (uint32_t)J;
        slots[sl] = sl;
    }
    __global__ void k_edge_heads(int nslots, const unsigned long long *skeys, const int *sslots, int *first_flag)
    {
        const int p = blockIdx.x * blockDim.x + threadIdx.x;
        if (p >= nslots)
            return;
        if (p == 0 || skeys[p] != skeys[p - 1])
            first_flag[sslots[p]] = 1; // stable sort: the first slot of the run is the smallest one
    }
    __global__ void k_edge_records(int nslots, const unsigned long long *skeys, const int *sslots, const int *first_rank, const int *prims, int *eot, int *sil_rec)
    {
        const int p = blockIdx.x * blockDim.x + threadIdx.x;
        if (p >= nslots || (p > 0 && skeys[p] == skeys[p - 1]))
            return;
        const int e = first_rank[sslots[p]];
        int *rec = sil_rec + (size_t)e * 4;
        int r0 = -1, r1 = -1, r2 = -1, r3 = -1;
        for (int q = p; q < nslots && skeys[q] == skeys[p]; ++q)
        {
            const int sl = sslots[q];
            eot[sl] = e;
            const int i = sl / 3, j = sl - 3 * i;
            const int I = j - 1 < 0 ? 2 : j - 1;
            int J = j, K = j + 1 > 2 ? 0 : j + 1;
            int orientation = 1;
            if (prims[3 * i + J] > prims[3 * i + K])
            {
                const int t = J;
                J = K;
                K = t;
                orientation = -1;
            }
            if (orientation == 1)
                r0 = prims[3 * i + I];
            else
                r3 = prims[3 * i + I];
            r1 = prims[3 * i + J];
            r2 = prims[3 * i + K];
        }
        rec[0] = r0, rec[1] = r1, rec[2] = r2, rec[3] = r3;
    }
    __global__ void k_prim_edges(int nprims, const int *first_flag, const int *eot, int *prim_sil)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= nprims)
            return;
        int out[3] = {-1, -1, -1};
        int ptr = 0;
        for (int j = 0; j < 3; ++j)
            if (first_flag[3 * i + j])
                out[ptr++] = eot[3 * i + j];
        prim_sil[3 * i] = out[0], prim_sil[3 * i + 1] = out[1], prim_sil[3 * i + 2] = out[2];
    }
    // fills s->d_sil_rec (E x 4), s->d_prim_sil (nprims x 3), s->nsil = E from s->d_prims
    static wost_status device_silhouettes_3d(wost_scene *s)
    {
        const int n = s->nprims, ns = 3 * n, B = 256;
        cudaStream_t st = 0;
        DevBuf<unsigned long long> keys, skeys;
        DevBuf<int> slots, sslots, first_flag, first_rank, eot;
        DevBuf<char> temp;
        WOST_CUDA_TRY(keys.alloc(ns));
        WOST_CUDA_TRY(skeys.alloc(ns));
        WOST_CUDA_TRY(slots.alloc(ns));
        WOST_CUDA_TRY(sslots.alloc(ns));
        WOST_CUDA_TRY(first_flag.alloc(ns));
        WOST_CUDA_TRY(first_rank.alloc(ns));
        WOST_CUDA_TRY(eot.alloc(ns));
        k_edge_keys<<<div_up(ns, B), B, 0, st>>>(n, s->d_prims.ptr, keys.ptr, slots.ptr);
        size_t tb1 = 0, tb2 = 0;
        WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tb1, keys.ptr, skeys.ptr, slots.ptr, sslots.ptr, ns, 0, 64, st));
        WOST_CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tb2, first_flag.ptr, first_rank.ptr, ns, st));
        WOST_CUDA_TRY(temp.alloc(std::max(tb1, tb2)));
        WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(temp.ptr, tb1, keys.ptr, skeys.ptr, slots.ptr, sslots.ptr, ns, 0, 64, st));
        WOST_CUDA_TRY(cudaMemsetAsync(first_flag.ptr, 0, sizeof(int) * (size_t)ns, st));
        k_edge_heads<<<div_up(ns, B), B, 0, st>>>(ns, skeys.ptr, sslots.ptr, first_flag.ptr);
        WOST_CUDA_TRY(cub::DeviceScan::ExclusiveSum(temp.ptr, tb2, first_flag.ptr, first_rank.ptr, ns, st));
        int last_rank = 0, last_flag = 0;
        WOST_CUDA_TRY(cudaMemcpyAsync(&last_rank, first_rank.ptr + (ns - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(&last_flag, first_flag.ptr + (ns - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        const int E = last_rank + last_flag;
        s->nsil = E;
        WOST_CUDA_TRY(s->d_sil_rec.alloc((size_t)E * 4));
        WOST_CUDA_TRY(s->d_prim_sil.alloc((size_t)n * 3));
        k_edge_records<<<div_up(ns, B), B, 0, st>>>(ns, skeys.ptr, sslots.ptr, first_rank.ptr, s->d_prims.ptr, eot.ptr, s->d_sil_rec.ptr);
        k_prim_edges<<<div_up(n, B), B, 0, st>>>(n, first_flag.ptr, eot.ptr, s->d_prim_sil.ptr);
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }

    template <int D>
    static wost_status build_scene(wost_scene *s)
    {
        const int n = s->nprims;
        const int NN = 2 * n - 1;
        const int B = 256;
        cudaStream_t st = 0;
        DevBuf<float> lboxes, lcones, partial;
        DevBuf<uint32_t> code_in, idx_in;
        DevBuf<int> flags, dup, leaf_count, leaf_first;
        DevBuf<float> nsum, obb;
        DevBuf<float> ext;
        DevBuf<unsigned long long> key64;
        DevBuf<char> temp;
        WOST_CUDA_TRY(lboxes.alloc((size_t)n * 2 * D));
        WOST_CUDA_TRY(lcones.alloc((size_t)n * (D + 2)));
        WOST_CUDA_TRY(code_in.alloc(n));
        WOST_CUDA_TRY(idx_in.alloc(n));
        WOST_CUDA_TRY(s->d_morton.alloc(n));
        WOST_CUDA_TRY(s->d_order.alloc(n));
        WOST_CUDA_TRY(s->d_nodes.alloc(NN));
        WOST_CUDA_TRY(s->d_boxes.alloc((size_t)NN * 2 * D));
        WOST_CUDA_TRY(s->d_cones.alloc((size_t)NN * (D + 2)));
        WOST_CUDA_TRY(flags.alloc(std::max(1, n - 1)));
        WOST_CUDA_TRY(leaf_count.alloc(std::max(1, n - 1)));
        WOST_CUDA_TRY(leaf_first.alloc(std::max(1, n - 1)));
        WOST_CUDA_TRY(ext.alloc((size_t)NN * 2 * D));
        WOST_CUDA_TRY(s->d_leaf_rank.alloc(n));
        WOST_CUDA_TRY(dup.alloc(1));
        WOST_CUDA_TRY(cudaMemsetAsync(dup.ptr, 0, sizeof(int), st));

        k_leaf_boxes_cones<D><<<div_up(n, B), B, 0, st>>>(n, s->d_verts.ptr, s->d_prims.ptr, s->d_sil_rec.ptr, s->d_prim_sil.ptr, lboxes.ptr, lcones.ptr);
        const int rb = std::min(1024, div_up(n, 256));
        WOST_CUDA_TRY(partial.alloc((size_t)(rb + 1) * 2 * D));
        k_reduce_boxes<D><<<rb, 256, 0, st>>>(n, lboxes.ptr, partial.ptr);
        k_reduce_boxes<D><<<1, 256, 0, st>>>(rb, partial.ptr, partial.ptr + (size_t)rb * 2 * D);
        const float *whole = partial.ptr + (size_t)rb * 2 * D;
        k_morton<D><<<div_up(n, B), B, 0, st>>>(n, lboxes.ptr, whole, code_in.ptr, idx_in.ptr);
        size_t temp_bytes = 0;
        WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, code_in.ptr, s->d_morton.ptr, idx_in.ptr, s->d_order.ptr, n, 0, 32, st));
        WOST_CUDA_TRY(temp.alloc(temp_bytes));
        WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(temp.ptr, temp_bytes, code_in.ptr, s->d_morton.ptr, idx_in.ptr, s->d_order.ptr, n, 0, 32, st));
        if (n > 1)
            k_check_unique<<<div_up(n - 1, B), B, 0, st>>>(n, s->d_morton.ptr, dup.ptr);
        int h_dup = 0;
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_dup, dup.ptr, sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        s->used_morton64 = h_dup != 0;

        WOST_CUDA_TRY(cudaMemsetAsync(s->d_nodes.ptr, 0xFF, sizeof(Node) * (size_t)NN, st));
        k_init_leaves<D><<<div_up(n, B), B, 0, st>>>(n, s->d_order.ptr, lboxes.ptr, lcones.ptr, s->d_nodes.ptr, s->d_boxes.ptr, s->d_cones.ptr);
        if (n > 1)
        {
            if (s->used_morton64)
            {
                WOST_CUDA_TRY(key64.alloc(n));
                k_keys64<<<div_up(n, B), B, 0, st>>>(n, s->d_morton.ptr, s->d_order.ptr, key64.ptr);
                k_karras<unsigned long long><<<div_up(n - 1, B), B, 0, st>>>(n, key64.ptr, s->d_nodes.ptr, leaf_count.ptr, leaf_first.ptr);
            }
            else
                k_karras<uint32_t><<<div_up(n - 1, B), B, 0, st>>>(n, s->d_morton.ptr, s->d_nodes.ptr, leaf_count.ptr, leaf_first.ptr);
            WOST_CUDA_TRY(cudaMemsetAsync(flags.ptr, 0, sizeof(int) * (size_t)(n - 1), st));
            k_refit_boxes<D><<<div_up(n, B), B, 0, st>>>(n, s->d_nodes.ptr, s->d_boxes.ptr, flags.ptr);
            WOST_CUDA_TRY(cudaMemsetAsync(flags.ptr, 0, sizeof(int) * (size_t)(n - 1), st));
            k_refit_cones<D><<<div_up(n, B), B, 0, st>>>(n, s->d_nodes.ptr, s->d_boxes.ptr, s->d_cones.ptr, flags.ptr);
        }
        k_leaf_extents<D><<<div_up(n, B), B, 0, st>>>(n, s->d_order.ptr, s->d_verts.ptr, s->d_prims.ptr, ext.ptr);
        if (n > 1)
        {
            WOST_CUDA_TRY(cudaMemsetAsync(flags.ptr, 0, sizeof(int) * (size_t)(n - 1), st));
            k_refit_boxes<D><<<div_up(n, B), B, 0, st>>>(n, s->d_nodes.ptr, ext.ptr, flags.ptr);
        }
        k_leaf_rank<<<div_up(n, B), B, 0, st>>>(n, s->d_nodes.ptr, leaf_count.ptr, s->d_leaf_rank.ptr);
        // traversal layout
        WOST_CUDA_TRY(s->d_tnodes.alloc(sizeof(TNode<D>) * (size_t)std::max(1, n - 1)));
        WOST_CUDA_TRY(s->d_cnodes.alloc(sizeof(CNode<D>) * (size_t)std::max(1, n - 1)));
        WOST_CUDA_TRY(s->d_leaf_prim.alloc(sizeof(LeafPrim<D>) * (size_t)n));
        WOST_CUDA_TRY(s->d_leaf_sil.alloc(sizeof(LeafSil<D>) * (size_t)n * D));
        WOST_CUDA_TRY(s->d_leaf_obj.alloc(n));
        k_layout_leaves<D><<<div_up(n, B), B, 0, st>>>(n, s->d_order.ptr, s->d_verts.ptr, s->d_prims.ptr, s->d_sil_rec.ptr, s->d_prim_sil.ptr,
                                                       reinterpret_cast<LeafPrim<D> *>(s->d_leaf_prim.ptr), s->d_leaf_obj.ptr,
                                                       reinterpret_cast<LeafSil<D> *>(s->d_leaf_sil.ptr));
        if (n > 1)
        {
            // oriented boxes of every node (pruning bound of the closest-point query, lbvh.cuh: Obb)
            const LeafPrim<D> *lp = reinterpret_cast<const LeafPrim<D> *>(s->d_leaf_prim.ptr);
            WOST_CUDA_TRY(nsum.alloc((size_t)NN * D));
            WOST_CUDA_TRY(obb.alloc((size_t)NN * obb_floats<D>()));
            WOST_CUDA_TRY(cudaMemsetAsync(flags.ptr, 0, sizeof(int) * (size_t)(n - 1), st));
            k_obb_normals<D><<<div_up(n, B), B, 0, st>>>(n, s->d_nodes.ptr, lp, nsum.ptr, flags.ptr);
            constexpr int kSmall = 8, kMedium = 1024;
            k_obb_extents<D, 1><<<div_up(NN, 256), 256, 0, st>>>(n, leaf_first.ptr, leaf_count.ptr, lp, nsum.ptr, ext.ptr, obb.ptr, 0, kSmall);
            k_obb_extents<D, 32><<<div_up(n - 1, 256 / 32), 256, 0, st>>>(n, leaf_first.ptr, leaf_count.ptr, lp, nsum.ptr, ext.ptr, obb.ptr, kSmall, kMedium);
            k_obb_extents<D, 256><<<n - 1, 256, 0, st>>>(n, leaf_first.ptr, leaf_count.ptr, lp, nsum.ptr, ext.ptr, obb.ptr, kMedium, 0x7fffffff);
            // record order: breadth-first for the top levels (as many levels as fit the L2 window budget), Karras order below
            DevBuf<int> depth, hist;
            DevBuf<unsigned long long> bkey_in, bkey_out;
            DevBuf<uint32_t> bval_in, bval_out, perm;
            WOST_CUDA_TRY(depth.alloc(n - 1));
            WOST_CUDA_TRY(hist.alloc(64));
            WOST_CUDA_TRY(bkey_in.alloc(n - 1));
            WOST_CUDA_TRY(bkey_out.alloc(n - 1));
            WOST_CUDA_TRY(bval_in.alloc(n - 1));
            WOST_CUDA_TRY(bval_out.alloc(n - 1));
            WOST_CUDA_TRY(perm.alloc(n - 1));
            WOST_CUDA_TRY(cudaMemsetAsync(hist.ptr, 0, 64 * sizeof(int), st));
            k_node_depth<<<div_up(n - 1, B), B, 0, st>>>(n, s->d_nodes.ptr, depth.ptr, hist.ptr);
            int h_hist[64];
            WOST_CUDA_TRY(cudaMemcpyAsync(h_hist, hist.ptr, sizeof(h_hist), cudaMemcpyDeviceToHost, st));
            WOST_CUDA_TRY(cudaStreamSynchronize(st));
            int top_depth = 0;
            size_t top_nodes = 0;
            while (top_depth < 63 && (top_nodes + (size_t)h_hist[top_depth]) * sizeof(TNode<D>) <= kL2WindowBudget)
                top_nodes += (size_t)h_hist[top_depth++];
            s->top_nodes = (int)top_nodes;
            s->top_depth = top_depth;
            k_bfs_keys<<<div_up(n - 1, B), B, 0, st>>>(n, depth.ptr, top_depth, bkey_in.ptr, bval_in.ptr);
            size_t tb2 = 0;
            WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tb2, bkey_in.ptr, bkey_out.ptr, bval_in.ptr, bval_out.ptr, n - 1, 0, 40, st));
            DevBuf<char> temp2;
            WOST_CUDA_TRY(temp2.alloc(tb2));
            WOST_CUDA_TRY(cub::DeviceRadixSort::SortPairs(temp2.ptr, tb2, bkey_in.ptr, bkey_out.ptr, bval_in.ptr, bval_out.ptr, n - 1, 0, 40, st));
            k_perm_scatter<<<div_up(n - 1, B), B, 0, st>>>(n, bval_out.ptr, perm.ptr);
            k_layout_internal<D><<<div_up(n - 1, B), B, 0, st>>>(n, s->d_nodes.ptr, ext.ptr, s->d_cones.ptr, obb.ptr, perm.ptr,
                                                                  reinterpret_cast<TNode<D> *>(s->d_tnodes.ptr), reinterpret_cast<CNode<D> *>(s->d_cnodes.ptr));
        }
        WOST_CUDA_TRY(cudaMemcpyAsync(s->root_box, s->d_boxes.ptr, sizeof(float) * 2 * D, cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(s->root_cone, s->d_cones.ptr, sizeof(float) * (D + 2), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(s->root_ext, ext.ptr, sizeof(float) * 2 * D, cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(s->d_overflow.alloc(32));
        WOST_CUDA_TRY(cudaMemsetAsync(s->d_overflow.ptr, 0, 32 * sizeof(int), st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }

    // ------------------------------------------------------------------ query kernels (parity probes + SDF channels)
    constexpr int kQueryBlock = 128;

    template <int D>
    __global__ void __launch_bounds__(kQueryBlock) k_nearest(DevBVH<D> bvh, const float *pts, int n, uint32_t *idx, float *dist, int *overflow)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, overflow, spill_mem);
        // warp-uniform trip count: query_nearest is warp-synchronous (lanes past the end pass valid = false)
        for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x)
        {
            const int i = base + threadIdx.x;
            const bool valid = i < n;
            Vec<D> q;
#pragma unroll
            for (int k = 0; k < D; ++k)
                q[k] = valid ? pts[(size_t)i * D + k] : 0.0f;
            uint32_t pos = kInvalid;
            float d = 0.0f;
            query_nearest(bvh, q, st, pos, d, valid);
            if (valid)
            {
                if (idx)
                    idx[i] = (pos == kInvalid) ? kInvalid : __ldg(bvh.leaf_obj + pos);
                dist[i] = d;
            }
        }
    }
    template <int D>
    __global__ void __launch_bounds__(kQueryBlock) k_silhouette(DevBVH<D> bvh, const float *pts, int n, int flip, float *dist, int *overflow)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, overflow, spill_mem);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        {
            Vec<D> q;
#pragma unroll
            for (int k = 0; k < D; ++k)
                q[k] = pts[(size_t)i * D + k];
            dist[i] = query_silhouette(bvh, q, flip != 0, st);
        }
    }
    template <int D, bool TestOnly>
    __global__ void __launch_bounds__(kQueryBlock) k_ray(DevBVH<D> bvh, const float *org, const float *dir, const float *max_dist, int n,
                                                         int *hit, float *t, float *uv, uint32_t *obj, int *overflow)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, overflow, spill_mem);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        {
            Vec<D> o, d;
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                o[k] = org[(size_t)i * D + k];
                d[k] = dir[(size_t)i * D + k];
            }
            const RayResult r = query_ray<D, TestOnly>(bvh, Ray<D>(o, d), max_dist[i], st);
            hit[i] = r.hit ? 1 : 0;
            if (!TestOnly)
            {
                t[i] = r.t;
                uv[2 * (size_t)i] = r.hit ? r.u : 0.0f;
                uv[2 * (size_t)i + 1] = r.hit ? r.v : 0.0f;
                obj[i] = (r.leaf_pos == kInvalid) ? kInvalid : __ldg(bvh.leaf_obj + r.leaf_pos);
            }
        }
    }
    template <int D>
    __global__ void k_sample_sphere(DevBVH<D> bvh, const float *ctr, const float *radius, const float *u, int n, int *obj, float *pdf)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        Vec<D> c;
#pragma unroll
        for (int k = 0; k < D; ++k)
            c[k] = ctr[(size_t)i * D + k];
        int pos;
        float p;
        sample_in_sphere(bvh, c, radius[i], u[i], pos, p);
        obj[i] = pos < 0 ? -1 : (int)__ldg(bvh.leaf_obj + pos);
        pdf[i] = p;
    }

    static int query_grid(int n)
    {
        const int blocks = div_up(n, kQueryBlock);
        const int cap = sm_count() * 8;
        return std::max(1, std::min(blocks, cap));
    }
    wost_status scene_traversal_stats(const wost_scene *s, uint64_t *full, uint64_t *restricted, int reset)
    {
        int h[4] = {0, 0, 0, 0};
        WOST_CUDA_TRY(cudaMemcpy(h, s->d_overflow.ptr, sizeof(h), cudaMemcpyDeviceToHost));
        if (full)
            *full = (uint64_t)(unsigned)h[1];
        if (restricted)
            *restricted = (uint64_t)(unsigned)h[2];
        if (reset)
            WOST_CUDA_TRY(cudaMemset(s->d_overflow.ptr + 1, 0, 2 * sizeof(int)));
        return WOST_OK;
    }
    static wost_status check_overflow(const wost_scene *s)
    {
        int h = 0;
        WOST_CUDA_TRY(cudaMemcpy(&h, s->d_overflow.ptr, sizeof(int), cudaMemcpyDeviceToHost));
        if (h)
        {
            cudaMemset(s->d_overflow.ptr, 0, sizeof(int));
            set_last_error("LBVH traversal stack overflow (tree deeper than 64 levels)");
            return WOST_ERR_OVERFLOW;
        }
        return WOST_OK;
    }
    // ------------------------------------------------------------------ closest-point queries without any allocation
    // Staging area of wost_lbvh_nearest_static(): a fixed array in the library's device image, so the entry point never asks
    // the driver (or libwost's own block cache) for memory. Every GPU that loads libwost.so carries these ~1.25 MiB until the
    // process exits. Layout in 32-bit words:
    //   [0, 64)                         control words of the traversal (SmemStack::overflow: [0] stack-overflow flag,
    //                                   [1], [2] traversal statistics, [8, 24) debug counters) -- zeroed at the start of each call
    //   [64, 64 + 3 M)                  query points (n * dim floats)
    //   [64 + 3 M, 64 + 4 M)            closest primitive per query
    //   [64 + 4 M, 64 + 5 M)            distance per query                       with M = kStaticNearestMaxPoints
    constexpr int kStaticNearestMaxPoints = WOST_NEAREST_STATIC_MAX_POINTS;
    constexpr int kStaticCtrlWords = 64;
    constexpr size_t kStaticScratchWords = kStaticCtrlWords + (size_t)5 * kStaticNearestMaxPoints;
    static __device__ __align__(256) uint32_t g_nearest_static_scratch[kStaticScratchWords];

    // In host code `g_nearest_static_scratch` is only a handle for the runtime, not somewhere a kernel can read or write: the
    // device address has to be asked for, per device, while that device is current. It is looked up once per GPU and kept.
    constexpr int kStaticMaxDevices = 64;
    static std::mutex g_static_scratch_mu;
    static uint32_t *g_static_scratch_addr[kStaticMaxDevices] = {};
    static wost_status static_scratch_for_current_device(int device, uint32_t **out)
    {
        std::lock_guard<std::mutex> lock(g_static_scratch_mu);
        if (!g_static_scratch_addr[device])
        {
            uint32_t *p = nullptr;
            WOST_CUDA_TRY(cudaGetSymbolAddress((void **)&p, g_nearest_static_scratch));
            if (!p)
            {
                set_last_error("wost_lbvh_nearest_static: cudaGetSymbolAddress returned a null scratch address");
                return WOST_ERR_CUDA;
            }
            g_static_scratch_addr[device] = p;
        }
        *out = g_static_scratch_addr[device];
        return WOST_OK;
    }
} // namespace wost

using namespace wost;

extern "C"
{
    const char *wost_last_error(void) { return wost::last_error().c_str(); }
    const char *wost_version(void) { return "wost-b200 0.1 (sm_100a)"; }
    int wost_device_count(void)
    {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess)
            return 0;
        return n;
    }
    wost_status wost_trim_allocation_cache(void)
    {
        AllocCache &c = alloc_cache();
        std::lock_guard<std::mutex> lock(c.mu);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto &kv : c.free_blocks)
        {
            cudaSetDevice((int)(kv.first >> 48));
            cudaFree(kv.second);
        }
        cudaSetDevice(cur);
        c.free_blocks.clear();
        c.cached_bytes = 0;
        return WOST_OK;
    }
    wost_status wost_set_device(int device)
    {
        WOST_CUDA_TRY(cudaSetDevice(device));
        return WOST_OK;
    }

    wost_status wost_scene_create(int dim, const float *verts, int nverts, const int *prims, int nprims, wost_scene **out)
    {
        if (!out || !verts || !prims || (dim != 2 && dim != 3) || nverts <= 0 || nprims <= 0)
        {
            set_last_error("wost_scene_create: invalid argument");
            return WOST_ERR_INVALID;
        }
        for (size_t i = 0; i < (size_t)nprims * dim; ++i)
            if (prims[i] < 0 || prims[i] >= nverts)
            {
                set_last_error("wost_scene_create: primitive index out of range");
                return WOST_ERR_INVALID;
            }
        wost_scene *s = new wost_scene();
        s->dim = dim;
        cudaGetDevice(&s->device);
        s->nverts = nverts;
        s->nprims = nprims;
        s->h_verts.assign(verts, verts + (size_t)nverts * dim);
        s->h_prims.assign(prims, prims + (size_t)nprims * dim);
        // silhouette adjacency: vertices of a polyline on the host (two linear passes), edges of a triangle mesh on the device
        // (WOST_HOST_SILHOUETTES=1 keeps the host hash-map path as a cross-check)
        std::vector<int> sil_rec, prim_sil;
        const bool host_sil = dim == 2 || getenv("WOST_HOST_SILHOUETTES") != nullptr;
        if (dim == 2)
            host_silhouettes_2d(nverts, nprims, prims, sil_rec, prim_sil);
        else if (host_sil)
            host_silhouettes_3d(nprims, prims, sil_rec, prim_sil);
        s->nsil = (int)sil_rec.size() / 4;
        wost_status rc = WOST_OK;
        auto fail = [&](wost_status r)
        {
            delete s;
            return r;
        };
        if (s->d_verts.upload(verts, (size_t)nverts * dim) != cudaSuccess || s->d_prims.upload(prims, (size_t)nprims * dim) != cudaSuccess ||
            (host_sil && (s->d_sil_rec.upload(sil_rec.data(), sil_rec.size()) != cudaSuccess || s->d_prim_sil.upload(prim_sil.data(), prim_sil.size()) != cudaSuccess)))
        {
            set_last_error(std::string("wost_scene_create: upload failed: ") + cudaGetErrorString(cudaGetLastError()));
            return fail(WOST_ERR_CUDA);
        }
        if (!host_sil)
        {
            rc = device_silhouettes_3d(s);
            if (rc != WOST_OK)
                return fail(rc);
        }
        rc = (dim == 2) ? build_scene<2>(s) : build_scene<3>(s);
        if (rc != WOST_OK)
            return fail(rc);
        *out = s;
        return WOST_OK;
    }
    void wost_scene_destroy(wost_scene *scene) { delete scene; }

    wost_status wost_scene_info(const wost_scene *s, int *dim, int *num_prims, int *num_nodes, int *used_morton64)
    {
        if (!s)
            return WOST_ERR_INVALID;
        if (dim)
            *dim = s->dim;
        if (num_prims)
            *num_prims = s->nprims;
        if (num_nodes)
            *num_nodes = 2 * s->nprims - 1;
        if (used_morton64)
            *used_morton64 = s->used_morton64 ? 1 : 0;
        return WOST_OK;
    }
    wost_status wost_scene_dump_bvh(const wost_scene *s, uint32_t *nodes, float *aabbs, float *cones)
    {
        if (!s)
            return WOST_ERR_INVALID;
        const size_t NN = (size_t)2 * s->nprims - 1;
        if (nodes)
            WOST_CUDA_TRY(cudaMemcpy(nodes, s->d_nodes.ptr, NN * sizeof(Node), cudaMemcpyDeviceToHost));
        if (aabbs)
            WOST_CUDA_TRY(cudaMemcpy(aabbs, s->d_boxes.ptr, NN * 2 * s->dim * sizeof(float), cudaMemcpyDeviceToHost));
        if (cones)
            WOST_CUDA_TRY(cudaMemcpy(cones, s->d_cones.ptr, NN * (s->dim + 2) * sizeof(float), cudaMemcpyDeviceToHost));
        return WOST_OK;
    }
    wost_status wost_scene_dump_traversal_layout(const wost_scene *s, float *records, int *record_floats, float *leaf_verts)
    {
        if (!s)
            return WOST_ERR_INVALID;
        const size_t rec = s->dim == 2 ? sizeof(TNode<2>) : sizeof(TNode<3>);
        const size_t leaf = s->dim == 2 ? sizeof(LeafPrim<2>) : sizeof(LeafPrim<3>);
        if (record_floats)
            *record_floats = (int)(rec / sizeof(float));
        if (records && s->nprims > 1)
            WOST_CUDA_TRY(cudaMemcpy(records, s->d_tnodes.ptr, rec * (size_t)(s->nprims - 1), cudaMemcpyDeviceToHost));
        if (leaf_verts && s->nprims > 0)
        {
            std::vector<char> raw(leaf * (size_t)s->nprims);
            WOST_CUDA_TRY(cudaMemcpy(raw.data(), s->d_leaf_prim.ptr, raw.size(), cudaMemcpyDeviceToHost));
            const int nf = s->dim * s->dim;
            for (int i = 0; i < s->nprims; ++i)
                memcpy(leaf_verts + (size_t)i * nf, raw.data() + (size_t)i * leaf, sizeof(float) * nf);
        }
        return WOST_OK;
    }
    wost_status wost_scene_traversal_stats(const wost_scene *s, uint64_t *full_reference_order, uint64_t *restricted_second_pass, int reset)
    {
        if (!s)
            return WOST_ERR_INVALID;
        return scene_traversal_stats(s, full_reference_order, restricted_second_pass, reset);
    }
    wost_status wost_scene_dump_sorted(const wost_scene *s, uint32_t *morton, uint32_t *object_idx)
    {
        if (!s)
            return WOST_ERR_INVALID;
        if (morton)
            WOST_CUDA_TRY(cudaMemcpy(morton, s->d_morton.ptr, (size_t)s->nprims * 4, cudaMemcpyDeviceToHost));
        if (object_idx)
            WOST_CUDA_TRY(cudaMemcpy(object_idx, s->d_order.ptr, (size_t)s->nprims * 4, cudaMemcpyDeviceToHost));
        return WOST_OK;
    }

    // ---- device-pointer variants
    wost_status wost_lbvh_nearest_dev(const wost_scene *s, const float *d_pts, int n, uint32_t *d_idx, float *d_dist, void *stream)
    {
        if (!s || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        cudaStream_t st = (cudaStream_t)stream;
        const size_t sm = stack_smem_bytes(kQueryBlock);
        if (s->dim == 2)
            k_nearest<2><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<2>(), d_pts, n, d_idx, d_dist, s->d_overflow.ptr);
        else
            k_nearest<3><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<3>(), d_pts, n, d_idx, d_dist, s->d_overflow.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }
    wost_status wost_lbvh_silhouette_dev(const wost_scene *s, const float *d_pts, int n, int flip, float *d_dist, void *stream)
    {
        if (!s || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        cudaStream_t st = (cudaStream_t)stream;
        const size_t sm = stack_smem_bytes(kQueryBlock);
        if (s->dim == 2)
            k_silhouette<2><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<2>(), d_pts, n, flip, d_dist, s->d_overflow.ptr);
        else
            k_silhouette<3><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<3>(), d_pts, n, flip, d_dist, s->d_overflow.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }
    wost_status wost_lbvh_ray_dev(const wost_scene *s, const float *d_org, const float *d_dir, const float *d_max_dist, int n,
                                  int *d_hit, float *d_t, float *d_uv, uint32_t *d_obj, void *stream)
    {
        if (!s || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        cudaStream_t st = (cudaStream_t)stream;
        const size_t sm = stack_smem_bytes(kQueryBlock);
        if (s->dim == 2)
            k_ray<2, false><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<2>(), d_org, d_dir, d_max_dist, n, d_hit, d_t, d_uv, d_obj, s->d_overflow.ptr);
        else
            k_ray<3, false><<<query_grid(n), kQueryBlock, sm, st>>>(s->view<3>(), d_org, d_dir, d_max_dist, n, d_hit, d_t, d_uv, d_obj, s->d_overflow.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }

    // ---- host-buffer variants
    wost_status wost_lbvh_nearest(const wost_scene *s, const float *pts, int n, uint32_t *idx, float *dist)
    {
        if (!s || !pts || !dist || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        DevBuf<float> d_pts, d_dist;
        DevBuf<uint32_t> d_idx;
        WOST_CUDA_TRY(d_pts.upload(pts, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_dist.alloc(n));
        WOST_CUDA_TRY(d_idx.alloc(n));
        wost_status rc = wost_lbvh_nearest_dev(s, d_pts.ptr, n, d_idx.ptr, d_dist.ptr, nullptr);
        if (rc != WOST_OK)
            return rc;
        WOST_CUDA_TRY(d_dist.download(dist, n));
        if (idx)
            WOST_CUDA_TRY(d_idx.download(idx, n));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return check_overflow(s);
    }
    wost_status wost_lbvh_silhouette(const wost_scene *s, const float *pts, int n, int flip, float *dist)
    {
        if (!s || !pts || !dist || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        DevBuf<float> d_pts, d_dist;
        WOST_CUDA_TRY(d_pts.upload(pts, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_dist.alloc(n));
        wost_status rc = wost_lbvh_silhouette_dev(s, d_pts.ptr, n, flip, d_dist.ptr, nullptr);
        if (rc != WOST_OK)
            return rc;
        WOST_CUDA_TRY(d_dist.download(dist, n));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return check_overflow(s);
    }
    wost_status wost_lbvh_ray(const wost_scene *s, const float *org, const float *dir, const float *max_dist, int n,
                              int *hit, float *t, float *uv, uint32_t *obj)
    {
        if (!s || !org || !dir || !max_dist || !hit || !t || !uv || !obj || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        DevBuf<float> d_org, d_dir, d_md, d_t, d_uv;
        DevBuf<int> d_hit;
        DevBuf<uint32_t> d_obj;
        WOST_CUDA_TRY(d_org.upload(org, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_dir.upload(dir, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_md.upload(max_dist, n));
        WOST_CUDA_TRY(d_t.alloc(n));
        WOST_CUDA_TRY(d_uv.alloc((size_t)n * 2));
        WOST_CUDA_TRY(d_hit.alloc(n));
        WOST_CUDA_TRY(d_obj.alloc(n));
        wost_status rc = wost_lbvh_ray_dev(s, d_org.ptr, d_dir.ptr, d_md.ptr, n, d_hit.ptr, d_t.ptr, d_uv.ptr, d_obj.ptr, nullptr);
        if (rc != WOST_OK)
            return rc;
        WOST_CUDA_TRY(d_hit.download(hit, n));
        WOST_CUDA_TRY(d_t.download(t, n));
        WOST_CUDA_TRY(d_uv.download(uv, (size_t)n * 2));
        WOST_CUDA_TRY(d_obj.download(obj, n));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return check_overflow(s);
    }
    wost_status wost_lbvh_ray_any(const wost_scene *s, const float *org, const float *dir, const float *max_dist, int n, int *hit)
    {
        if (!s || !org || !dir || !max_dist || !hit || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        DevBuf<float> d_org, d_dir, d_md;
        DevBuf<int> d_hit;
        WOST_CUDA_TRY(d_org.upload(org, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_dir.upload(dir, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_md.upload(max_dist, n));
        WOST_CUDA_TRY(d_hit.alloc(n));
        const size_t sm = stack_smem_bytes(kQueryBlock);
        if (s->dim == 2)
            k_ray<2, true><<<query_grid(n), kQueryBlock, sm>>>(s->view<2>(), d_org.ptr, d_dir.ptr, d_md.ptr, n, d_hit.ptr, nullptr, nullptr, nullptr, s->d_overflow.ptr);
        else
            k_ray<3, true><<<query_grid(n), kQueryBlock, sm>>>(s->view<3>(), d_org.ptr, d_dir.ptr, d_md.ptr, n, d_hit.ptr, nullptr, nullptr, nullptr, s->d_overflow.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(d_hit.download(hit, n));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return check_overflow(s);
    }
    wost_status wost_lbvh_sample_sphere(const wost_scene *s, const float *ctr, const float *radius, const float *u, int n, int *obj, float *pdf)
    {
        if (!s || !ctr || !radius || !u || !obj || !pdf || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        DevBuf<float> d_c, d_r, d_u, d_pdf;
        DevBuf<int> d_obj;
        WOST_CUDA_TRY(d_c.upload(ctr, (size_t)n * s->dim));
        WOST_CUDA_TRY(d_r.upload(radius, n));
        WOST_CUDA_TRY(d_u.upload(u, n));
        WOST_CUDA_TRY(d_pdf.alloc(n));
        WOST_CUDA_TRY(d_obj.alloc(n));
        if (s->dim == 2)
            k_sample_sphere<2><<<div_up(n, 128), 128>>>(s->view<2>(), d_c.ptr, d_r.ptr, d_u.ptr, n, d_obj.ptr, d_pdf.ptr);
        else
            k_sample_sphere<3><<<div_up(n, 128), 128>>>(s->view<3>(), d_c.ptr, d_r.ptr, d_u.ptr, n, d_obj.ptr, d_pdf.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(d_obj.download(obj, n));
        WOST_CUDA_TRY(d_pdf.download(pdf, n));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return WOST_OK;
    }

    // ---- host-buffer closest-point query that allocates nothing (include/wost.h has the contract)
    wost_status wost_lbvh_nearest_static(const wost_scene *s, const float *pts, int n, uint32_t *idx, float *dist)
    {
        if (!s || !pts || !dist || n < 0)
        {
            set_last_error("wost_lbvh_nearest_static: invalid argument");
            return WOST_ERR_INVALID;
        }
        if (n > kStaticNearestMaxPoints)
        {
            char buf[192];
            snprintf(buf, sizeof(buf), "wost_lbvh_nearest_static: %d query points exceed the fixed scratch of %d points; split the batch or call wost_lbvh_nearest",
                     n, kStaticNearestMaxPoints);
            set_last_error(buf);
            return WOST_ERR_INVALID;
        }
        if (n == 0)
            return WOST_OK;
        int cur = -1;
        WOST_CUDA_TRY(cudaGetDevice(&cur));
        if (cur != s->device || cur < 0 || cur >= kStaticMaxDevices)
        {
            char buf[192];
            snprintf(buf, sizeof(buf), "wost_lbvh_nearest_static: the scene lives on device %d but device %d is current (wost_set_device first; devices 0..%d)",
                     s->device, cur, kStaticMaxDevices - 1);
            set_last_error(buf);
            return WOST_ERR_INVALID;
        }
        uint32_t *scratch = nullptr;
        wost_status rc = static_scratch_for_current_device(cur, &scratch);
        if (rc != WOST_OK)
            return rc;
        int *d_ctrl = reinterpret_cast<int *>(scratch);
        float *d_pts = reinterpret_cast<float *>(scratch + kStaticCtrlWords);
        uint32_t *d_idx = scratch + kStaticCtrlWords + (size_t)3 * kStaticNearestMaxPoints;
        float *d_dist = reinterpret_cast<float *>(scratch + kStaticCtrlWords + (size_t)4 * kStaticNearestMaxPoints);
        // the array still holds whatever the previous call left: the overflow flag and the counters must start at zero
        WOST_CUDA_TRY(cudaMemsetAsync(d_ctrl, 0, kStaticCtrlWords * sizeof(uint32_t), nullptr));
        WOST_CUDA_TRY(cudaMemcpyAsync(d_pts, pts, (size_t)n * s->dim * sizeof(float), cudaMemcpyHostToDevice, nullptr));
        const size_t sm = stack_smem_bytes(kQueryBlock);
        if (s->dim == 2)
            k_nearest<2><<<query_grid(n), kQueryBlock, sm>>>(s->view<2>(), d_pts, n, d_idx, d_dist, d_ctrl);
        else
            k_nearest<3><<<query_grid(n), kQueryBlock, sm>>>(s->view<3>(), d_pts, n, d_idx, d_dist, d_ctrl);
        WOST_CUDA_TRY(cudaGetLastError());
        int overflowed = 0;
        WOST_CUDA_TRY(cudaMemcpyAsync(dist, d_dist, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, nullptr));
        if (idx)
            WOST_CUDA_TRY(cudaMemcpyAsync(idx, d_idx, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, nullptr));
        WOST_CUDA_TRY(cudaMemcpyAsync(&overflowed, d_ctrl, sizeof(int), cudaMemcpyDeviceToHost, nullptr));
        WOST_CUDA_TRY(cudaStreamSynchronize(nullptr));
        if (overflowed)
        {
            set_last_error("LBVH traversal stack overflow (tree deeper than 64 levels)");
            return WOST_ERR_OVERFLOW;
        }
        return WOST_OK;
    }
}
