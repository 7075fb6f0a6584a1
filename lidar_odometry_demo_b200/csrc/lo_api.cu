// Context, error reporting, scratch buffers, voxel-key dump and parameter handling of liblo_b200.
#include <cstring>
#include <fstream>
#include <sstream>

#include "lo_host.h"

namespace lo {

static thread_local std::string g_last_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

int DevBuf::ensure(size_t bytes) {
    if (bytes <= cap) return LO_OK;
    size_t want = cap ? cap : 4096;
    while (want < bytes) want *= 2;
    if (p) LO_CUDA(cudaFree(p));
    p = nullptr;
    cap = 0;
    LO_CUDA(cudaMalloc(&p, want));
    cap = want;
    return LO_OK;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}
int PinnedBuf::ensure(size_t bytes) {
    if (bytes <= cap) return LO_OK;
    size_t want = cap ? cap : 4096;
    while (want < bytes) want *= 2;
    if (p) LO_CUDA(cudaFreeHost(p));
    p = nullptr;
    cap = 0;
    LO_CUDA(cudaMallocHost(&p, want));
    cap = want;
    return LO_OK;
}
void PinnedBuf::release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
}

__global__ void voxel_keys_kernel(const PointXYZ* __restrict__ pts, uint32_t n, float voxel_size,
                                  long long* __restrict__ keys3, uint32_t* __restrict__ hash22) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PointXYZ p = pts[i];
    const long long ix = axis_index(p.x, voxel_size), iy = axis_index(p.y, voxel_size), iz = axis_index(p.z, voxel_size);
    if (keys3) {
        keys3[3 * i] = ix;
        keys3[3 * i + 1] = iy;
        keys3[3 * i + 2] = iz;
    }
    if (hash22) hash22[i] = reference_hash22(ix, iy, iz);
}

}  // namespace lo

using namespace lo;

extern "C" const char* lo_last_error(void) { return g_last_error.c_str(); }

extern "C" int lo_ctx_create(int device, lo_ctx** out) {
    if (!out) return LO_E_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        set_error("no CUDA device available (%s); liblo_b200 has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return LO_E_CUDA;
    }
    if (device < 0 || device >= count) {
        set_error("device %d out of range (have %d)", device, count);
        return LO_E_ARG;
    }
    LO_CUDA(cudaSetDevice(device));
    lo_ctx* ctx = new lo_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    LO_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    LO_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    {   // grids allocate stream-ordered (lo_grid.cu); keep what they free in the pool instead of returning it to the OS
        cudaMemPool_t pool;
        LO_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        uint64_t keep = ~0ull;
        LO_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    if (ctx->small.ensure(64 * 1024) != LO_OK) return LO_E_CUDA;
    *out = ctx;
    return LO_OK;
}

extern "C" void lo_ctx_destroy(lo_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (DevBuf* b : {&ctx->in_a, &ctx->in_b, &ctx->out_a, &ctx->out_b, &ctx->tmp_a, &ctx->tmp_b, &ctx->tmp_c, &ctx->tmp_d, &ctx->small})
        b->release();
    ctx->pin_in.release();
    ctx->pin_out.release();
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" void* lo_ctx_stream(lo_ctx* ctx) { return ctx ? static_cast<void*>(ctx->stream) : nullptr; }

extern "C" int lo_ctx_synchronize(lo_ctx* ctx) {
    if (!ctx) return LO_E_ARG;
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

extern "C" uint64_t lo_ctx_kernel_launches(const lo_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int lo_voxel_keys(lo_ctx* ctx, float voxel_size, const float* xyz16, size_t n, int64_t* keys3, uint32_t* hash22) {
    if (!ctx || (!xyz16 && n)) return LO_E_ARG;
    if (n == 0) return LO_OK;
    LO_CUDA(cudaSetDevice(ctx->device));
    LO_TRY(ctx->in_a.ensure(n * sizeof(PointXYZ)));
    LO_TRY(ctx->out_a.ensure(n * 3 * sizeof(long long)));
    LO_TRY(ctx->out_b.ensure(n * sizeof(uint32_t)));
    LO_CUDA(cudaMemcpyAsync(ctx->in_a.p, xyz16, n * sizeof(PointXYZ), cudaMemcpyHostToDevice, ctx->stream));
    voxel_keys_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(ctx->in_a.as<PointXYZ>(), static_cast<uint32_t>(n), voxel_size,
                                                                   ctx->out_a.as<long long>(), ctx->out_b.as<uint32_t>());
    LO_LAUNCHED(ctx);
    if (keys3) LO_CUDA(cudaMemcpyAsync(keys3, ctx->out_a.p, n * 3 * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    if (hash22) LO_CUDA(cudaMemcpyAsync(hash22, ctx->out_b.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    LO_CUDA(cudaStreamSynchronize(ctx->stream));
    return LO_OK;
}

// LidarOdometry::Params::GetROSDeclaration defaults (src/lidar_odometry.h:36-48)
extern "C" void lo_params_default(lo_params* p) {
    if (!p) return;
    p->lidar_min_range = 4.0f;
    p->lidar_max_range = 80.0f;
    p->keyframe_voxel_size = 0.2f;
    p->keyframe_max_points_cnt = 20;
    p->keyframe_matching_voxel_size = 0.3f;
    p->keyframe_update_voxel_size = 0.1f;
    p->keyframe_cleanup_range = 80.0f;
    p->angular_divergence_threshold = 5.0f;
}

// Minimal reader for the reference's config/params.yaml: picks `name: value` lines of the eight
// lidar_odometry.* parameters wherever they appear (the file nests them under /**: ros__parameters:
// lidar_odometry:). Unknown keys are ignored; missing keys keep their defaults.
extern "C" int lo_params_from_yaml(const char* path, lo_params* p) {
    if (!path || !p) return LO_E_ARG;
    lo_params_default(p);
    std::ifstream f(path);
    if (!f) {
        set_error("cannot open %s", path);
        return LO_E_IO;
    }
    std::string line;
    while (std::getline(f, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.resize(hash);
        const size_t colon = line.find(':');
        if (colon == std::string::npos) continue;
        std::string key = line.substr(0, colon), val = line.substr(colon + 1);
        auto trim = [](std::string& s) {
            const size_t a = s.find_first_not_of(" \t\r\n\"");
            const size_t b = s.find_last_not_of(" \t\r\n\"");
            s = (a == std::string::npos) ? std::string() : s.substr(a, b - a + 1);
        };
        trim(key);
        trim(val);
        if (val.empty()) continue;
        char* end = nullptr;
        const double v = std::strtod(val.c_str(), &end);
        if (end == val.c_str()) continue;
        if (key == "lidar_min_range") p->lidar_min_range = static_cast<float>(v);
        else if (key == "lidar_max_range") p->lidar_max_range = static_cast<float>(v);
        else if (key == "keyframe_voxel_size") p->keyframe_voxel_size = static_cast<float>(v);
        else if (key == "keyframe_max_points_cnt") p->keyframe_max_points_cnt = static_cast<uint64_t>(v);
        else if (key == "keyframe_matching_voxel_size") p->keyframe_matching_voxel_size = static_cast<float>(v);
        else if (key == "keyframe_update_voxel_size") p->keyframe_update_voxel_size = static_cast<float>(v);
        else if (key == "keyframe_cleanup_range") p->keyframe_cleanup_range = static_cast<float>(v);
        else if (key == "angular_divergence_threshold") p->angular_divergence_threshold = static_cast<float>(v);
    }
    return LO_OK;
}
