// laser_slam_b200/csrc/zhsm_handoff.cu -- the two hand-over formats either side of the matchers.
//
// 1. covariance -> information: what the pose graph consumes from a matcher result.
//    CMapNode::matchNodePairICP (src/mapGraph/MapNode.cpp:638-651): the six numbers of CICPWarpper::getResult become a
//    symmetric 3x3 (cov[0], cov[3] times `scale` = 1), information = covariance.inverse() (Eigen); the loop-closure
//    caller then REPLACES it by the constant matrix m_cov_loopnode = diag(10,10,10) when m_bUseConstantCov
//    (src/mapGraph/MapGraph.cpp:250-261, 1427-1429: setInformation(m_cov_loopnode) -- the "covariance" constant is used
//    as the information matrix as it is, reproduced).  One thread per record.
//    Eigen's fixed-size 3x3 inverse is the cofactor formula (Eigen/src/LU/InverseImpl.h, compute_inverse<.., 3>):
//    inverse(i,j) = cofactor(j,i) / det, det = c00*m00 + (c10*m10 + c20*m20).  Eigen is not in this image: the sum
//    order of det follows Eigen 3's unrolled redux ("PARITY UNPINNED" to the last bit, exact in the formula).
//
// 2. the client -> server frame record: (x, y, theta, syn_num, [cov[6]], np, r[np]) as
//    CClientFrontend::sendFrameInfo[Cov] emits it (src/tcp_slam/clientFrontend.cpp:300-305, 405-407), written by
//    CClientLocal (src/tcp_slam/clientLocal.cpp:189-199) and parsed by CServerSocket::readClient
//    (src/tcp_slam/serverSocket.cpp:60-115) through a QDataStream: big-endian, a quint16 block length first, and --
//    QDataStream's default since Qt 4.6 -- every float written as an 8-byte double (ZHSM_FRAME_QT_SINGLE selects the
//    4-byte form of QDataStream::SinglePrecision / Qt < 4.6).  Host code: packing bytes is not GPU work; the scans land
//    in a scan-pool row ready for zhsm_*_batch.  Qt is not in this image: the layout follows QDataStream's documented
//    serialisation, checked by a hand-built byte vector and round trips.
#include <math.h>
#include <string.h>

#include "zhsm_internal.h"

namespace {

__device__ __forceinline__ double cof(const double m[9], int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m[3 * i1 + j1] * m[3 * i2 + j2] - m[3 * i1 + j2] * m[3 * i2 + j1];
}

__global__ void cov_to_information_kernel(const double *cov6, int M, double scale, int use_const, double c0, double c1,
                                          double c2, double *info, unsigned char *ok) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    const double *c = cov6 + 6 * (size_t)k;
    // MapNode.cpp:640-647: covariance.fill(0); (0,0) = cov[0]*scale; (1,1) = cov[3]*scale; the rest as they are
    const double m[9] = {c[0] * scale, c[1], c[2], c[1], c[3] * scale, c[4], c[2], c[4], c[5]};
    const double c00 = cof(m, 0, 0), c10 = cof(m, 1, 0), c20 = cof(m, 2, 0);
    const double det = c00 * m[0] + (c10 * m[3] + c20 * m[6]);
    const double invdet = 1.0 / det;
    double *o = info + 9 * (size_t)k;
    if (use_const) {  // MapGraph.cpp:1427-1429
        for (int q = 0; q < 9; q++) o[q] = 0.0;
        o[0] = c0; o[4] = c1; o[8] = c2;
    } else {
        o[0] = c00 * invdet; o[1] = c10 * invdet; o[2] = c20 * invdet;  // result.row(0) = cofactors_col0 * invdet
        o[3] = cof(m, 0, 1) * invdet; o[4] = cof(m, 1, 1) * invdet; o[5] = cof(m, 2, 1) * invdet;
        o[6] = cof(m, 0, 2) * invdet; o[7] = cof(m, 1, 2) * invdet; o[8] = cof(m, 2, 2) * invdet;
    }
    if (ok) ok[k] = (det != 0.0 && !isnan(det) && !isinf(invdet)) ? 1 : 0;
}

// ---- big-endian writers / readers
struct Wr {
    uint8_t *p, *end;
    bool over;
    void bytes(const void *v, int n) {
        if (p + n > end) { over = true; return; }
        const uint8_t *s = (const uint8_t *)v;
        for (int i = 0; i < n; i++) p[i] = s[n - 1 - i];  // host is little-endian (x86-64 / aarch64)
        p += n;
    }
    void u16(uint16_t v) { bytes(&v, 2); }
    void i32(int32_t v) { bytes(&v, 4); }
    void f64(double v) { bytes(&v, 8); }
    void real(float v, bool single) { if (single) bytes(&v, 4); else f64((double)v); }
};
struct Rd {
    const uint8_t *p, *end;
    bool over;
    void bytes(void *v, int n) {
        if (p + n > end) { over = true; memset(v, 0, n); return; }
        uint8_t *d = (uint8_t *)v;
        for (int i = 0; i < n; i++) d[n - 1 - i] = p[i];
        p += n;
    }
    uint16_t u16() { uint16_t v; bytes(&v, 2); return v; }
    int32_t i32() { int32_t v; bytes(&v, 4); return v; }
    double f64() { double v; bytes(&v, 8); return v; }
    float real(bool single) { if (single) { float v; bytes(&v, 4); return v; } return (float)f64(); }
};

}  // namespace

extern "C" zhsm_status zhsm_cov_to_information(zhsm_handle *h, const double *cov6, int M, double scale,
                                               const double *const_diag, double *info, uint8_t *ok) {
    if (!h) return ZHSM_ERR_ARG;
    if (!cov6 || !info || M < 0) { h->err = "zhsm_cov_to_information: bad argument"; return ZHSM_ERR_ARG; }
    if (M == 0) return ZHSM_OK;
    ZHSM_CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    StreamScope scope(h, st);
    void *p;
    zhsm_status s;
    if ((s = zhsm_stage(h, 17, sizeof(double) * 6 * (size_t)M, &p)) != ZHSM_OK) return s;
    double *d_cov = (double *)p;
    if ((s = zhsm_stage(h, 18, sizeof(double) * 9 * (size_t)M, &p)) != ZHSM_OK) return s;
    double *d_info = (double *)p;
    if ((s = zhsm_stage(h, 19, (size_t)M, &p)) != ZHSM_OK) return s;
    unsigned char *d_ok = (unsigned char *)p;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(d_cov, cov6, sizeof(double) * 6 * (size_t)M, cudaMemcpyHostToDevice, st));
    cov_to_information_kernel<<<(M + 127) / 128, 128, 0, st>>>(d_cov, M, scale, const_diag != nullptr,
                                                               const_diag ? const_diag[0] : 0, const_diag ? const_diag[1] : 0,
                                                               const_diag ? const_diag[2] : 0, d_info, d_ok);
    ZHSM_CUDA_TRY(h, cudaGetLastError());
    h->launches++;
    ZHSM_CUDA_TRY(h, cudaMemcpyAsync(info, d_info, sizeof(double) * 9 * (size_t)M, cudaMemcpyDeviceToHost, st));
    if (ok) ZHSM_CUDA_TRY(h, cudaMemcpyAsync(ok, d_ok, (size_t)M, cudaMemcpyDeviceToHost, st));
    ZHSM_CUDA_TRY(h, cudaStreamSynchronize(st));
    return ZHSM_OK;
}

extern "C" long long zhsm_frame_pack(const zhsm_frame *f, const float *r, unsigned flags, uint8_t *buf, long long cap) {
    if (!f || !buf || cap < 2 || f->np < 0 || f->np > 65535 || (f->np > 0 && !r)) return -1;
    const bool single = flags & ZHSM_FRAME_QT_SINGLE, with_cov = flags & ZHSM_FRAME_WITH_COV;
    Wr w = {buf, buf + cap, false};
    w.u16(0);  // block length, patched below (clientLocal.cpp:191,198-199)
    w.real(f->x, single); w.real(f->y, single); w.real(f->th, single);
    w.i32(f->syn);
    if (with_cov) for (int k = 0; k < 6; k++) w.f64(f->cov[k]);
    w.u16((uint16_t)f->np);
    for (int j = 0; j < f->np; j++) w.real(r[j], single);
    if (w.over) return -1;
    const long long total = w.p - buf;
    if (total - 2 > 65535) return -1;  // the reference's quint16 length field cannot say it
    Wr lw = {buf, buf + 2, false};
    lw.u16((uint16_t)(total - 2));
    return total;
}

extern "C" long long zhsm_frame_unpack(const uint8_t *buf, long long len, unsigned flags, zhsm_frame *f, float *r, int cap_r) {
    if (!f || len < 0) return -1;
    const bool single = flags & ZHSM_FRAME_QT_SINGLE, with_cov = flags & ZHSM_FRAME_WITH_COV;
    if (len < 2) return 0;  // "bytesAvailable() < sizeof(quint16)": wait for more (serverSocket.cpp:69-70)
    if (!buf) return -1;
    Rd rd = {buf, buf + len, false};
    const long long block = rd.u16();
    if (len < 2 + block) return 0;  // not enough data yet (:73-74)
    rd.end = buf + 2 + block;
    memset(f, 0, sizeof(*f));
    f->x = rd.real(single); f->y = rd.real(single); f->th = rd.real(single);
    f->syn = rd.i32();
    if (with_cov) for (int k = 0; k < 6; k++) f->cov[k] = rd.f64();
    f->np = rd.u16();
    if (rd.over || f->np > cap_r || (f->np > 0 && !r)) return -1;
    for (int j = 0; j < f->np; j++) r[j] = rd.real(single);
    if (rd.over) return -1;
    return 2 + block;  // bytes consumed; the reference trusts the length field the same way (nextBlockSize)
}
