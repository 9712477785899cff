// The loop of the reference's tool/extract_and_match_example.cc:57-101 (minus OpenCV image IO / imshow),
// written against include/sift_cuda/ exactly as it would be against the reference's headers.
//   g++ -std=c++17 -I include -I /usr/local/cuda/include examples/extract_and_match.cc \
//       -L another_cuda_sift_b200 -lb200sift -L /usr/local/cuda/lib64 -lcudart -o extract_and_match
#include <cstdio>
#include <random>
#include <vector>

#include "sift_cuda/interface/Detector.hh"
#include "sift_cuda/sift_func/Match.cuh"

int main(int argc, char** argv) {
    const int cols = argc > 1 ? std::atoi(argv[1]) : 752, rows = argc > 2 ? std::atoi(argv[2]) : 480;
    const int n_frames = argc > 3 ? std::atoi(argv[3]) : 3;
    std::vector<Imagef> host_img;
    std::mt19937 rng(1234);
    for (int f = 0; f < n_frames; f++) {  // blocky random texture, shifted by f pixels per frame
        Imagef img(rows, cols);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) {
                const unsigned h = (unsigned(((x + f) / 7) * 73856093u) ^ unsigned((y / 5) * 19349663u)) % 256u;
                img.at(y, x) = float((h + ((x + f) * 3 + y * 5) % 32) % 256);
            }
        host_img.push_back(img);
    }
    CudaSiftConfig config;
    config.upscale = false;
    config.numFeatures = 2000;
    config.col_width = int(host_img.front().m_image_size.col);
    config.row_width = int(host_img.front().m_image_size.row);
    sift_cuda::Detector detector(config);
    if (!detector.gpuWarmUpAndAllocate()) return 1;
    cudaDeviceSynchronize();

    int prev_size = 0;
    std::vector<int> want_kpts, want_good;
    for (size_t idx = 0; idx < host_img.size(); idx++) {
        detector.detectAndCompute(host_img[idx]);
        detector.copyToHost(false);
        cudaDeviceSynchronize();
        const int curr_size = detector.total_size;
        int good = 0;
        if (idx != 0) {
            const auto matches_host = sift_cuda::matchBruteForce(detector.prev_descriptor, prev_size, detector.device_descriptor, curr_size);
            for (int m : matches_host) good += m >= 0;
        }
        std::printf("frame %zu: %d keypoints, %d matches to the previous frame; first kpt (%.0f, %.0f)\n", idx, curr_size, good,
                    curr_size ? detector.final_kpts[0].x : 0.f, curr_size ? detector.final_kpts[0].y : 0.f);
        prev_size = curr_size;
        want_kpts.push_back(curr_size);
        want_good.push_back(good);
    }

    // the reference's public stage methods (Detector.hh:36-43), here compatibility shims over the latest frame
    detector.getGaussianPyramid();
    detector.getKeyPoints();
    detector.adjustKpts();
    detector.adjustOrientationAndCleanup();
    detector.extractDescriptor();
    detector.collectAcrossOctaves();
    if (detector.total_size != want_kpts.back()) {
        std::fprintf(stderr, "stage shims: %d keypoints, expected %d\n", detector.total_size, want_kpts.back());
        return 2;
    }

    // the same frames through the pipelined batch call on a fresh detector (same slot history), engine options spelled out
    b200sift_options opt;
    b200sift_default_options(&opt);
    opt.frames_in_flight = 4;
    sift_cuda::Detector batch_detector(config, opt);
    if (!batch_detector.gpuWarmUpAndAllocate()) return 1;
    const auto res = batch_detector.detectAndComputeBatch(host_img, true, true);
    for (size_t idx = 0; idx < res.size(); idx++) {
        int good = 0;
        for (int m : res[idx].matches_from_previous) good += m >= 0;
        std::printf("batch frame %zu: %d keypoints, %d matches to the previous frame\n", idx, res[idx].total_size, good);
        if (res[idx].total_size != want_kpts[idx] || good != want_good[idx]) {
            std::fprintf(stderr, "batch frame %zu differs from the blocking loop\n", idx);
            return 3;
        }
    }
    return 0;
}
