#include <cstdio>
#include "../../mc-slam_b200/csrc/match_kernels.cuh"
__global__ void k(const double* Rt, double x0, double y0, double x1, double y1) {
    const double* Pm[8] = {Rt, Rt + 12};
    double xs[8] = {x0, x1}, ys[8] = {y0, y1};
    double X[3];
    mcs::dlt_nviews<8>(Pm, xs, ys, 2, false, X);
    printf("4x4: %g %g %g\n", X[0], X[1], X[2]);
    mcs::dlt_nviews<8>(Pm, xs, ys, 2, true, X);
    printf("design: %g %g %g\n", X[0], X[1], X[2]);
    double A[16], V[16], z[4];
    for (int i = 0; i < 4; i++) {
        A[0 * 4 + i] = xs[0] * Pm[0][8 + i] - Pm[0][i];
        A[1 * 4 + i] = ys[0] * Pm[0][8 + i] - Pm[0][4 + i];
        A[2 * 4 + i] = xs[1] * Pm[1][8 + i] - Pm[1][i];
        A[3 * 4 + i] = ys[1] * Pm[1][8 + i] - Pm[1][4 + i];
    }
    for (int i = 0; i < 16; i++) printf("A[%d]=%g\n", i, A[i]);
    mcs::null_vector<double, 4, 4>(A, V, 4, 4, z, 2.220446049250313e-16);
    printf("z: %g %g %g %g\n", z[0], z[1], z[2], z[3]);
    for (int i = 0; i < 16; i++) printf("A'[%d]=%g V=%g\n", i, A[i], V[i]);
}
int main() {
    double Rt[24] = {1,0,0,0, 0,1,0,0, 0,0,1,0,   1,0,0,-0.165, 0,1,0,0, 0,0,1,0};
    double X[3] = {0.3, -0.2, 6.0};
    double *d; cudaMalloc(&d, sizeof(Rt)); cudaMemcpy(d, Rt, sizeof(Rt), cudaMemcpyHostToDevice);
    float u0=(float)(256*X[0]/X[2]+160), v0=(float)(256*X[1]/X[2]+120), u1=(float)(256*(X[0]-0.165)/X[2]+160);
    k<<<1,1>>>(d, ((double)u0-160)/256, ((double)v0-120)/256, ((double)u1-160)/256, ((double)v0-120)/256);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
