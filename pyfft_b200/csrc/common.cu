#include "common.cuh"

#include <math.h>
#include <stdarg.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

namespace pyfft {

static thread_local std::string g_last_error;

std::string& last_error_ref() { return g_last_error; }

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

int check_map(const pyfft_record_map* map, RecMap* out) {
    if (!map) return fail(PYFFT_EINVAL, "record map is NULL");
    if (map->nsig <= 0) return fail(PYFFT_EINVAL, "nsig must be positive (got %lld)", (long long)map->nsig);
    if (map->i0 < 0 || map->reflect < 0 || map->src_len <= 0 || map->src_base < 0)
        return fail(PYFFT_EINVAL, "record map has negative fields");
    if (map->reflect > 0 && map->i0 != 0)
        return fail(PYFFT_EINVAL, "reflect padding is only defined for i0 == 0 (fft_analysis.py:198)");
    out->nsig = map->nsig; out->i0 = map->i0; out->reflect = map->reflect;
    out->src_base = map->src_base; out->src_len = map->src_len;
    return PYFFT_OK;
}

int check_range(const RecMap& m, int64_t j_lo, int64_t j_hi) {
    if (j_lo < 0 || j_hi > m.nsig || j_lo > j_hi)
        return fail(PYFFT_EINVAL, "analysed index range [%lld, %lld) outside the record of %lld samples",
                    (long long)j_lo, (long long)j_hi, (long long)m.nsig);
    if (j_lo == j_hi) return PYFFT_OK;
    // the map is monotone on each side of j = reflect: check the end points of both pieces
    int64_t lo = INT64_MAX, hi = INT64_MIN;
    auto upd = [&](int64_t j) {
        int64_t e = src_index(m, j);
        if (e < lo) lo = e;
        if (e > hi) hi = e;
    };
    upd(j_lo); upd(j_hi - 1);
    if (j_lo < m.reflect && j_hi > m.reflect) { upd(m.reflect - 1); upd(m.reflect); }
    if (lo < 0 || hi >= m.src_len)
        return fail(PYFFT_EINVAL, "analysed range [%lld, %lld) needs source elements [%lld, %lld] but the "
                    "pointers cover [0, %lld)", (long long)j_lo, (long long)j_hi, (long long)lo,
                    (long long)hi, (long long)m.src_len);
    return PYFFT_OK;
}

int check_seg(const pyfft_segmentation* seg, int min_log2, int max_log2, int* log2l) {
    if (!seg) return fail(PYFFT_EINVAL, "segmentation is NULL");
    if (seg->nfft <= 0 || seg->hop <= 0 || seg->seg_first < 0 || seg->seg_count < 0)
        return fail(PYFFT_EINVAL, "segmentation needs nfft > 0, hop > 0, seg_first >= 0, seg_count >= 0");
    int l = ilog2(seg->nfft);
    if ((1 << l) != seg->nfft)
        return fail(PYFFT_EUNSUPPORTED, "nfft = %d is not a power of two", seg->nfft);
    if (l < min_log2 || l > max_log2)
        return fail(PYFFT_EUNSUPPORTED, "nfft = %d outside the supported range [%d, %d]", seg->nfft,
                    1 << min_log2, 1 << max_log2);
    *log2l = l;
    return PYFFT_OK;
}

static std::mutex g_tw_mutex;
static std::map<std::tuple<int, int, int64_t, int64_t, int64_t>, void*> g_tw_cache;

// table[k] = exp(-2 pi i k step / M), k = 0 .. count-1
const void* twiddle_table_ex(int dtype, int64_t M, int64_t step, int64_t count, int* status) {
    *status = PYFFT_OK;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { *status = fail(PYFFT_ECUDA, "cudaGetDevice failed"); return nullptr; }
    std::lock_guard<std::mutex> lock(g_tw_mutex);
    auto key = std::make_tuple(dev, dtype, M, step, count);
    auto it = g_tw_cache.find(key);
    if (it != g_tw_cache.end()) return it->second;
    size_t esz = dtype == PYFFT_F64 ? 16 : 8;
    std::vector<unsigned char> host(esz * (size_t)count);
    const long double two_pi = 6.283185307179586476925286766559005768L;
    for (int64_t k = 0; k < count; ++k) {
        // the angle is reduced exactly in integers before the long-double evaluation
        const int64_t j = (int64_t)(((__int128)k * step) % M);
        long double a = -two_pi * (long double)j / (long double)M;
        long double c = cosl(a), s = sinl(a);
        if (dtype == PYFFT_F64) {
            double* p = reinterpret_cast<double*>(host.data()) + 2 * k;
            p[0] = (double)c; p[1] = (double)s;
        } else {
            float* p = reinterpret_cast<float*>(host.data()) + 2 * k;
            p[0] = (float)c; p[1] = (float)s;
        }
    }
    void* d = nullptr;
    cudaError_t e = cudaMalloc(&d, host.size());
    if (e == cudaSuccess) e = cudaMemcpy(d, host.data(), host.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        *status = fail(PYFFT_ECUDA, "twiddle table upload failed: %s", cudaGetErrorString(e));
        if (d) cudaFree(d);
        return nullptr;
    }
    g_tw_cache[key] = d;
    return d;
}

const void* twiddle_table(int dtype, int64_t L, int* status) { return twiddle_table_ex(dtype, L, 1, L, status); }

void free_twiddles() {
    std::lock_guard<std::mutex> lock(g_tw_mutex);
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& kv : g_tw_cache) {
        cudaSetDevice(std::get<0>(kv.first));
        cudaFree(kv.second);
    }
    g_tw_cache.clear();
    cudaSetDevice(cur);
}

int device_sm_count() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}

}  // namespace pyfft

extern "C" {

const char* pyfft_version(void) { return "pyfft_b200 0.1.0 (sm_100a)"; }

const char* pyfft_last_error(void) { return pyfft::last_error_ref().c_str(); }

int pyfft_shutdown(void) {
    pyfft::free_twiddles();
    return PYFFT_OK;
}

int pyfft_sm_count(void) { return pyfft::device_sm_count(); }

}  // extern "C"
