#include <cstdio>
#include <vector>
#include "../../mc-slam_b200/csrc/match_kernels.cuh"
using namespace mcs;
int main() {
    const int C = 3, cap = 8, row_cap = 8;
    CalibDev cd[3];
    for (int c = 0; c < 3; c++) {
        double K[9] = {256,0,160, 0,256,120, 0,0,1};
        memcpy(cd[c].K, K, sizeof(K));
        double Rt[12] = {1,0,0,-0.165*c, 0,1,0,0, 0,0,1,0};
        memcpy(cd[c].Rt, Rt, sizeof(Rt));
        memcpy(cd[c].P, Rt, sizeof(Rt));
    }
    double X[3] = {0.3, -0.2, 6.0};
    std::vector<KeypointPod> kps(C * cap);
    for (int c = 0; c < 3; c++) {
        KeypointPod p{}; p.x = (float)(256 * (X[0] - 0.165 * c) / X[2] + 160); p.y = (float)(256 * X[1] / X[2] + 120);
        kps[c * cap + 1] = p;
    }
    int rows[row_cap * C] = {1, 1, -1,   1, 1, 1,   -1, 1, 1};
    int row_count = 3;
    std::vector<uint8_t> desc(C * cap * 32, 0);
    CalibDev* d_c; KeypointPod* d_k; int *d_r, *d_rc; uint8_t* d_d;
    cudaMalloc(&d_c, sizeof(cd)); cudaMemcpy(d_c, cd, sizeof(cd), cudaMemcpyHostToDevice);
    cudaMalloc(&d_k, kps.size() * sizeof(KeypointPod)); cudaMemcpy(d_k, kps.data(), kps.size() * sizeof(KeypointPod), cudaMemcpyHostToDevice);
    cudaMalloc(&d_r, sizeof(rows)); cudaMemcpy(d_r, rows, sizeof(rows), cudaMemcpyHostToDevice);
    cudaMalloc(&d_rc, 4); cudaMemcpy(d_rc, &row_count, 4, cudaMemcpyHostToDevice);
    cudaMalloc(&d_d, desc.size()); cudaMemcpy(d_d, desc.data(), desc.size(), cudaMemcpyHostToDevice);
    LandmarkOut lm;
    cudaMalloc(&lm.rows, row_cap * C * 4); cudaMalloc(&lm.accepted, row_cap); cudaMalloc(&lm.n_views, row_cap * 4);
    cudaMalloc(&lm.xyz, row_cap * 24); cudaMalloc(&lm.uv_ref, row_cap * 8); cudaMalloc(&lm.rep_cam, row_cap * 4); cudaMalloc(&lm.n_accepted, 4);
    cudaMemset(lm.n_accepted, 0, 4);
    triangulate_kernel<8><<<dim3(1, 1), 64>>>(C, cap, row_cap, 320, 240, d_c, d_r, d_rc, d_k, d_d, nullptr, lm);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    double xyz[row_cap * 3]; cudaMemcpy(xyz, lm.xyz, sizeof(xyz), cudaMemcpyDeviceToHost);
    for (int r = 0; r < 3; r++) printf("row %d: %g %g %g\n", r, xyz[r * 3], xyz[r * 3 + 1], xyz[r * 3 + 2]);
}
