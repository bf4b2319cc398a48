// plc_merge.cu — K8: rigid transform of each filtered scan into the map frame, colouring, and
// concatenation into the merged cloud.
//
// Replaces processRobotCloud + the concatenation in processPointClouds
// (/root/reference/src/pointcloud_merger.cpp:308-394): tf2::doTransform applies
// Eigen::Translation3f * Eigen::Quaternionf per point in float32, then every
// point becomes a pcl::PointXYZRGB with a constant colour and is appended to the merged cloud
// (sources in call order: robot first, then backhoe, :316-325) and to its own coloured cloud.
// One streaming kernel per source: N*step bytes read, N*32*2 bytes written.
#include "plc_internal.cuh"

namespace plc {
namespace {

struct Affine {
    float m[3][3];
    float t[3];
};

__device__ __forceinline__ float ld_f32m(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    unsigned b = (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
    return __uint_as_float(b);
}

// PointXYZRGB wire record (pcl::toROSMsg copies the 32-byte struct): x y z 1.0f | b g r a=255 | 12 zero bytes
__global__ void __launch_bounds__(256)
k_transform_colour(const unsigned char* __restrict__ data, unsigned n, unsigned step, unsigned ox, unsigned oy,
                   unsigned oz, bool aligned, Affine a, unsigned rgba, uint4* __restrict__ merged,
                   uint4* __restrict__ own) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned char* p = data + (size_t)i * step;
    float v0 = ld_f32m(p + ox, aligned), v1 = ld_f32m(p + oy, aligned), v2 = ld_f32m(p + oz, aligned);
    float r[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float acc = __fmul_rn(a.m[k][0], v0);
        acc = __fadd_rn(acc, __fmul_rn(a.m[k][1], v1));
        acc = __fadd_rn(acc, __fmul_rn(a.m[k][2], v2));
        r[k] = __fadd_rn(acc, a.t[k]);
    }
    uint4 lo = make_uint4(__float_as_uint(r[0]), __float_as_uint(r[1]), __float_as_uint(r[2]), 0x3f800000u);
    uint4 hi = make_uint4(rgba, 0u, 0u, 0u);
    merged[2 * (size_t)i] = lo;
    merged[2 * (size_t)i + 1] = hi;
    if (own) {
        own[2 * (size_t)i] = lo;
        own[2 * (size_t)i + 1] = hi;
    }
}

// Eigen::Quaternionf::toRotationMatrix in float32, translation narrowed to float
Affine make_affine(const double* t, const double* q) {
    const float qx = (float)q[0], qy = (float)q[1], qz = (float)q[2], qw = (float)q[3];
    const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
    const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    Affine a;
    a.m[0][0] = 1.0f - (tyy + tzz); a.m[0][1] = txy - twz;          a.m[0][2] = txz + twy;
    a.m[1][0] = txy + twz;          a.m[1][1] = 1.0f - (txx + tzz); a.m[1][2] = tyz - twx;
    a.m[2][0] = txz - twy;          a.m[2][1] = tyz + twx;          a.m[2][2] = 1.0f - (txx + tyy);
    a.t[0] = (float)t[0]; a.t[1] = (float)t[1]; a.t[2] = (float)t[2];
    return a;
}

}  // namespace
}  // namespace plc

using namespace plc;

extern "C" plc_status plc_merge(plc_ctx* ctx, const plc_merge_src* srcs, int n_src, void* out_merged,
                                uint64_t capacity, uint64_t* n_out, void* const* out_each, uint64_t* n_each) {
    PLC_RANGE();
    if (!ctx || n_src < 0 || (n_src && !srcs) || !n_out) return PLC_E_INVALID_ARG;
    *n_out = 0;
    uint64_t total = 0, raw_bytes = 0;
    for (int i = 0; i < n_src; ++i) {
        if (n_each) n_each[i] = 0;
        if (!srcs[i].present) continue;
        if (!cloud_view_ok(&srcs[i].cloud) || srcs[i].cloud.n_points > 0x7ffffff0ull) return PLC_E_INVALID_ARG;
        total += srcs[i].cloud.n_points;
        raw_bytes = raw_bytes > srcs[i].cloud.n_points * srcs[i].cloud.point_step
                        ? raw_bytes : srcs[i].cloud.n_points * srcs[i].cloud.point_step;
    }
    if (total > capacity || (total && !out_merged)) return PLC_E_CAPACITY;
    if (total == 0) return PLC_OK;
    PLC_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    PLC_CUDA_TRY(ctx, ctx->m_out.reserve(total * 32 * 2));      // [merged | per-source clouds back to back]
    uint4* merged = ctx->m_out.as<uint4>();
    uint4* each = merged + total * 2;
    uint64_t at = 0;
    ctx->timer.reset();
    ctx->times.merge_ms = 0.f;
    for (int i = 0; i < n_src; ++i) {
        const plc_merge_src& src = srcs[i];
        if (!src.present || src.cloud.n_points == 0) continue;
        const unsigned n = (unsigned)src.cloud.n_points;
        // the staging buffer is reused per source: order the copy after the previous kernel (same stream)
        const unsigned char* raw = nullptr;
        PLC_TRY(stage_bytes(ctx, src.cloud.data, (size_t)n * src.cloud.point_step, &raw));
        const bool aligned = ((src.cloud.point_step | src.cloud.off_x | src.cloud.off_y | src.cloud.off_z) & 3u) == 0;
        const unsigned rgba = 0xff000000u | ((unsigned)src.r << 16) | ((unsigned)src.g << 8) | (unsigned)src.b;
        const bool want_own = out_each && out_each[i];
        ctx->timer.begin(s);
        k_transform_colour<<<(n + 255) / 256, 256, 0, s>>>(raw, n, src.cloud.point_step,
                                                           src.cloud.off_x, src.cloud.off_y, src.cloud.off_z, aligned,
                                                           make_affine(src.t, src.q), rgba, merged + at * 2,
                                                           want_own ? each + at * 2 : nullptr);
        ctx->launches += 1;
        ctx->timer.end(s);
        if (want_own)
            PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_each[i], each + at * 2, (size_t)n * 32, cudaMemcpyDefault, s));
        if (n_each) n_each[i] = n;
        at += n;
    }
    PLC_CUDA_TRY(ctx, cudaGetLastError());
    PLC_CUDA_TRY(ctx, cudaMemcpyAsync(out_merged, merged, total * 32, cudaMemcpyDefault, s));
    PLC_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    ctx->times.merge_ms = ctx->timer.total_ms();
    *n_out = total;
    return PLC_OK;
}
