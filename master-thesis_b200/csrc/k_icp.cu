// k_icp.cu -- SURVEY 8(f) row n4: pcl::IterativeClosestPoint<PointXYZRGB, PointXYZRGB>::align as the reference configures it at
// cv_reconstruct/src/system/MappingSystem.cpp:321-341 (and the earlier copy :104-127): setMaxCorrespondenceDistance(5),
// setMaximumIterations(50), setTransformationEpsilon(1e-8), setEuclideanFitnessEpsilon(1); point-to-point, SVD (Umeyama without
// scaling), no reciprocal correspondences, no rejectors, identity guess.  PCL is absent from this image: the rules below follow
// PCL's registration/impl/icp.hpp and default_convergence_criteria.hpp as written down in DESIGN.md section 7 ("parity unpinned").
//
// Per iteration, on the device: nearest target point of every source point through the uniform grid the outlier filters use
// (expanding cubic shells with a proven stop; ties go to the lower point index, like a linear scan), and the 17 sums the
// transformation estimate needs {n, sum s, sum d, sum d s^T, sum |s - d|^2} in f64 with a FIXED reduction order (per-thread
// strided partial sums -> block tree -> one block over the block partials), so a run is deterministic.  On the host (C++): the
// 3x3 SVD, the f32 increment, the convergence rules.  The source cloud is transformed on the device by the f32 increment.
#include "cvr_internal.cuh"
#include <float.h>
#include <math.h>
#include <string.h>

#define ICP_THREADS 256
#define ICP_MAX_BLOCKS 1024
#define ICP_NSUM 17

// nearest point of the cell-sorted target cloud to q: squared distance (f32, x,y,z order, no FMA) and the target's original index
__device__ __forceinline__ void icp_nearest(const float4 q, const float4* __restrict__ sp, const int* __restrict__ cs, const SorGrid& g, float maxd,
                                            float& bestD, int& bestJ, float4& bestP)
{
    // the search starts from the cell of q clamped into the target's bounding box; `gap` is how far q lies outside it
    const float hx = g.mn[0] + (float)g.dim[0] * g.cell, hy = g.mn[1] + (float)g.dim[1] * g.cell, hz = g.mn[2] + (float)g.dim[2] * g.cell;
    const float qx = fminf(fmaxf(q.x, g.mn[0]), hx), qy = fminf(fmaxf(q.y, g.mn[1]), hy), qz = fminf(fmaxf(q.z, g.mn[2]), hz);
    const float gap = sqrtf((q.x - qx) * (q.x - qx) + (q.y - qy) * (q.y - qy) + (q.z - qz) * (q.z - qz));
    int cx, cy, cz;
    sor_cell(g, qx, qy, qz, cx, cy, cz);
    bestD = FLT_MAX; bestJ = 0x7fffffff;
    bestP = make_float4(0.f, 0.f, 0.f, 0.f);
    const int Rmax = max(max(g.dim[0], g.dim[1]), g.dim[2]);
    for (int R = 0; R <= Rmax; R++) {
        const int z0 = max(cz - R, 0), z1 = min(cz + R, g.dim[2] - 1);
        const int y0 = max(cy - R, 0), y1 = min(cy + R, g.dim[1] - 1);
        for (int iz = z0; iz <= z1; iz++) {
            const bool zf = (iz == cz - R) || (iz == cz + R);
            for (int iy = y0; iy <= y1; iy++) {
                const bool face = zf || (iy == cy - R) || (iy == cy + R);
                for (int part = 0; part < 2; part++) {                 // a whole x run on a face of the shell, else its two end caps
                    int xa, xb;
                    if (face) { if (part) break; xa = max(cx - R, 0); xb = min(cx + R, g.dim[0] - 1); }
                    else {
                        const int xx = part ? cx + R : cx - R;
                        if (part && R == 0) break;
                        if (xx < 0 || xx >= g.dim[0]) continue;
                        xa = xb = xx;
                    }
                    if (xa > xb) continue;
                    const int rowc = (iz * g.dim[1] + iy) * g.dim[0];
                    const int jb = cs[rowc + xa], je = cs[rowc + xb + 1];
                    for (int j = jb; j < je; j++) {
                        const float4 p = sp[j];
                        const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y), dz = __fsub_rn(q.z, p.z);
                        float d2 = __fmul_rn(dx, dx);
                        d2 = __fadd_rn(d2, __fmul_rn(dy, dy));
                        d2 = __fadd_rn(d2, __fmul_rn(dz, dz));
                        const int oj = __float_as_int(p.w);
                        if (d2 < bestD || (d2 == bestD && oj < bestJ)) { bestD = d2; bestJ = oj; bestP = p; }
                    }
                }
            }
        }
        // every unvisited target point is farther than R * cell from the clamped position, hence farther than R * cell - gap from q
        const float bound = (float)R * g.cell - gap;
        if (bound > 0.f && (bestD <= bound * bound * (1.0f - 1e-5f) || bound > maxd * (1.0f + 1e-5f))) break;
    }
}

// partial[block][17]: the sums over this block's source points whose nearest target point lies within maxd (maxd < 0: all points)
__global__ void __launch_bounds__(ICP_THREADS) k_icp_sums(const cvr_point* __restrict__ src, int ns, const float4* __restrict__ spts, const SorGrid* __restrict__ grid,
                                                          const int* __restrict__ cellstart, float maxd, double* __restrict__ partial)
{
    __shared__ double red[ICP_THREADS];
    const SorGrid g = grid[0];
    const float md2 = maxd < 0.f ? FLT_MAX : __fmul_rn(maxd, maxd);
    const float reach = maxd < 0.f ? FLT_MAX : maxd;
    double a[ICP_NSUM];
#pragma unroll
    for (int k = 0; k < ICP_NSUM; k++) a[k] = 0.0;
    const int chunk = (ns + gridDim.x - 1) / gridDim.x, beg = blockIdx.x * chunk, end = min(beg + chunk, ns);
    for (int i = beg + threadIdx.x; i < end; i += ICP_THREADS) {
        const float4 q = *(const float4*)(src + i);
        float d2; int j; float4 p;
        icp_nearest(q, spts, cellstart, g, reach, d2, j, p);
        if (j == 0x7fffffff || d2 > md2) continue;
        const double s[3] = { (double)q.x, (double)q.y, (double)q.z }, d[3] = { (double)p.x, (double)p.y, (double)p.z };
        a[0] += 1.0;
#pragma unroll
        for (int u = 0; u < 3; u++) {
            a[1 + u] += s[u]; a[4 + u] += d[u];
#pragma unroll
            for (int v = 0; v < 3; v++) a[7 + 3 * u + v] += d[u] * s[v];
        }
        a[16] += (double)d2;
    }
#pragma unroll 1
    for (int k = 0; k < ICP_NSUM; k++) {                               // fixed-order tree over the block
        red[threadIdx.x] = a[k];
        __syncthreads();
        for (int o = ICP_THREADS / 2; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
        if (threadIdx.x == 0) partial[(size_t)blockIdx.x * ICP_NSUM + k] = red[0];
        __syncthreads();
    }
}
__global__ void k_icp_final(const double* __restrict__ partial, int nblocks, double* __restrict__ out)
{
    const int k = threadIdx.x;
    if (k >= ICP_NSUM) return;
    double s = 0.0;
    for (int b = 0; b < nblocks; b++) s += partial[(size_t)b * ICP_NSUM + k];
    out[k] = s;
}

// ---- host side: Umeyama without scaling from the sums (f64, one-sided Jacobi SVD of the 3x3 covariance) ----
static void svd3(double A[3][3], double U[3][3], double S[3], double V[3][3])
{
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) V[i][j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0.0;
        for (int p = 0; p < 2; p++)
            for (int q = p + 1; q < 3; q++) {
                double al = 0, be = 0, ga = 0;
                for (int i = 0; i < 3; i++) { al += A[i][p] * A[i][p]; be += A[i][q] * A[i][q]; ga += A[i][p] * A[i][q]; }
                if (fabs(ga) <= 1e-300 || fabs(ga) <= 1e-17 * sqrt(al * be)) continue;
                off += fabs(ga);
                const double zeta = (be - al) / (2.0 * ga);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int i = 0; i < 3; i++) {
                    const double ap = A[i][p], aq = A[i][q];
                    A[i][p] = cs * ap - sn * aq; A[i][q] = sn * ap + cs * aq;
                    const double vp = V[i][p], vq = V[i][q];
                    V[i][p] = cs * vp - sn * vq; V[i][q] = sn * vp + cs * vq;
                }
            }
        if (off == 0.0) break;
    }
    for (int j = 0; j < 3; j++) S[j] = sqrt(A[0][j] * A[0][j] + A[1][j] * A[1][j] + A[2][j] * A[2][j]);
    int ord[3] = { 0, 1, 2 };
    for (int x = 0; x < 3; x++) for (int y = x + 1; y < 3; y++) if (S[ord[y]] > S[ord[x]]) { const int t = ord[x]; ord[x] = ord[y]; ord[y] = t; }
    const double tiny = 1e-12 * (S[ord[0]] > 0 ? S[ord[0]] : 1.0);
    for (int r = 0; r < 3; r++) {
        const int j = ord[r];
        if (S[j] > tiny) { for (int i = 0; i < 3; i++) U[i][j] = A[i][j] / S[j]; continue; }
        double u[3];
        if (r == 2) {                                                  // rank 2: the cross product of the two good columns
            const int x = ord[0], y = ord[1];
            u[0] = U[1][x] * U[2][y] - U[2][x] * U[1][y]; u[1] = U[2][x] * U[0][y] - U[0][x] * U[2][y]; u[2] = U[0][x] * U[1][y] - U[1][x] * U[0][y];
        } else if (r == 1) {                                           // rank 1: any unit vector orthogonal to the good column
            const int x = ord[0];
            const int m = fabs(U[0][x]) < fabs(U[1][x]) ? (fabs(U[0][x]) < fabs(U[2][x]) ? 0 : 2) : (fabs(U[1][x]) < fabs(U[2][x]) ? 1 : 2);
            double e3[3] = { 0, 0, 0 }; e3[m] = 1.0;
            for (int i = 0; i < 3; i++) u[i] = e3[i] - U[m][x] * U[i][x];
        } else { u[0] = 1; u[1] = 0; u[2] = 0; }
        const double nn = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        for (int i = 0; i < 3; i++) U[i][j] = u[i] / nn;
    }
}
static double det3(double M[3][3])
{
    return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) + M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
}
static void umeyama(const double* sums, float* T)
{
    const double n = sums[0];
    double ms[3], md[3], A[3][3], U[3][3], S[3], V[3][3];
    for (int i = 0; i < 3; i++) { ms[i] = sums[1 + i] / n; md[i] = sums[4 + i] / n; }
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) A[i][j] = sums[7 + 3 * i + j] / n - md[i] * ms[j];
    svd3(A, U, S, V);
    double sg[3] = { 1, 1, 1 };
    if (det3(U) * det3(V) < 0) { int m = 0; for (int j = 1; j < 3; j++) if (S[j] < S[m]) m = j; sg[m] = -1; }
    for (int i = 0; i < 3; i++) {
        double t = md[i];
        for (int j = 0; j < 3; j++) {
            double r = 0;
            for (int k = 0; k < 3; k++) r += U[i][k] * sg[k] * V[j][k];
            T[4 * i + j] = (float)r;
            t -= r * ms[j];
        }
        T[4 * i + 3] = (float)t;
    }
    T[12] = T[13] = T[14] = 0.f; T[15] = 1.f;
}
static void mat4_mul(const float* A, const float* B, float* C)         // C = A * B in f32 (Eigen::Matrix4f product)
{
    float R[16];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { float v = 0.f; for (int k = 0; k < 4; k++) v += A[4 * i + k] * B[4 * k + j]; R[4 * i + j] = v; }
    memcpy(C, R, sizeof(R));
}

extern "C" void cvr_icp_default_params(cvr_icp_params* p)
{
    if (!p) return;
    p->max_correspondence_distance = 5.0;    // MappingSystem.cpp:329
    p->max_iterations = 50;                  // :331
    p->transformation_epsilon = 1e-8;        // :333
    p->euclidean_fitness_epsilon = 1.0;      // :335
}

extern "C" int cvr_icp_align(cvr_ctx* c, int mem, const cvr_point* source, int ns, const cvr_point* target, int nt, const cvr_icp_params* prm,
                             cvr_point* aligned, cvr_icp_result* res)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (!source || !target || !prm || !res || ns < 1 || nt < 1 || (mem != CVR_MEM_HOST && mem != CVR_MEM_DEVICE)) { c->err = "cvr_icp_align: bad argument"; return CVR_ERR_INVALID_ARG; }
    if (ns > c->maxN || nt > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (!(prm->max_correspondence_distance > 0.0) || prm->max_iterations < 1) { c->err = "cvr_icp_align: bad parameters"; return CVR_ERR_INVALID_ARG; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    if (!c->d_icp) CVR_CUDA(c, cudaMalloc((void**)&c->d_icp, sizeof(double) * ICP_NSUM * (ICP_MAX_BLOCKS + 1)));
    if (!c->h_icp) CVR_CUDA(c, cudaHostAlloc((void**)&c->h_icp, sizeof(double) * ICP_NSUM, cudaHostAllocDefault));
    const cudaMemcpyKind kin = mem == CVR_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    const cudaMemcpyKind kout = mem == CVR_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    // target -> d_vox (sorted by cell into d_spts), source -> d_cloud (ping) / d_sorout (pong)
    CVR_CUDA(c, cudaMemcpyAsync(c->d_vox, target, (size_t)nt * sizeof(cvr_point), kin, c->stream));
    CVR_CUDA(c, cudaMemcpyAsync(c->d_cloud, source, (size_t)ns * sizeof(cvr_point), kin, c->stream));
    c->h_counts[1] = nt;
    CVR_CUDA(c, cudaMemcpyAsync(c->d_nvox, &c->h_counts[1], sizeof(int), cudaMemcpyHostToDevice, c->stream));
    SorGrid* grid; int* cellstart;
    int rc = cvr_build_point_grid(c, 1, c->d_vox, (size_t)c->maxN, c->d_nvox, false, 0.f, &grid, &cellstart);
    if (rc) return rc;
    const int nblocks = min(ICP_MAX_BLOCKS, cvr_div_up(ns, ICP_THREADS));
    double* partial = c->d_icp; double* total = c->d_icp + (size_t)ICP_MAX_BLOCKS * ICP_NSUM;
    auto sums = [&](const cvr_point* cur, float maxd) -> int {
        k_icp_sums<<<nblocks, ICP_THREADS, 0, c->stream>>>(cur, ns, c->d_spts, grid, cellstart, maxd, partial);
        LAUNCH_CHECK(c);
        k_icp_final<<<1, 32, 0, c->stream>>>(partial, nblocks, total);
        LAUNCH_CHECK(c);
        CVR_CUDA(c, cudaMemcpyAsync(c->h_icp, total, sizeof(double) * ICP_NSUM, cudaMemcpyDeviceToHost, c->stream));
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { c->err = std::string("stream sync: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
        return CVR_OK;
    };
    cvr_point* cur = c->d_cloud; cvr_point* nxt = c->d_sorout;
    float Tf[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
    double prev = DBL_MAX;
    int it = 0, state = 0, ncorr = 0;
    for (;;) {
        rc = sums(cur, (float)prm->max_correspondence_distance);
        if (rc) return rc;
        ncorr = (int)c->h_icp[0];
        if (ncorr < 3) { state = 0; break; }                           // PCL: "Not enough correspondences found"
        float Ti[16];
        umeyama(c->h_icp, Ti);
        rc = cvr_launch_transform(c, cur, nxt, ns, Ti);
        if (rc) return rc;
        { cvr_point* t = cur; cur = nxt; nxt = t; }
        mat4_mul(Ti, Tf, Tf);
        it++;
        if (it >= prm->max_iterations) { state = 1; break; }
        const double cosang = 0.5 * ((double)Ti[0] + (double)Ti[5] + (double)Ti[10] - 1.0);
        const double tsq = (double)Ti[3] * Ti[3] + (double)Ti[7] * Ti[7] + (double)Ti[11] * Ti[11];
        if (cosang >= 1.0 - prm->transformation_epsilon && tsq <= prm->transformation_epsilon) { state = 2; break; }
        const double mse = c->h_icp[16] / c->h_icp[0];
        if (fabs(mse - prev) < 1e-12) { state = 3; break; }
        if (fabs(mse - prev) / prev < prm->euclidean_fitness_epsilon) { state = 4; break; }
        prev = mse;
    }
    rc = sums(cur, -1.0f);                                             // getFitnessScore(): every aligned source point against its nearest target point
    if (rc) return rc;
    memcpy(res->T, Tf, sizeof(Tf));
    res->iterations = it; res->state = state; res->converged = state != 0; res->correspondences = ncorr;
    res->fitness = c->h_icp[0] > 0 ? c->h_icp[16] / c->h_icp[0] : 0.0;
    if (aligned) {
        CVR_CUDA(c, cudaMemcpyAsync(aligned, cur, (size_t)ns * sizeof(cvr_point), kout, c->stream));
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { c->err = std::string("stream sync: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
    }
    return CVR_OK;
}
