// Runs detectAndCompute on one image n times: the reference's tool/detection_example.cc:16-52 (profile it with
// ncu), with the same options -p/--path and -d/--debug_path, reading PGM instead of going through cv::imread.
//   g++ -std=c++17 -I include -I tools/cli -I /usr/local/cuda/include tools/cli/detection_example.cc \
//       -L another_cuda_sift_b200 -lb200sift -L /usr/local/cuda/lib64 -lcudart -o detection_example
#include <chrono>
#include <iostream>

#include "image_io.hh"
#include "sift_cuda/interface/Detector.hh"
#include "sift_cuda/types/CudaSiftConfig.hh"

int main(int argc, char** argv) {
    if (cli::flag(argc, argv, "-h", "--help")) {
        std::cout << "Runs detectAndCompute on one single image n times (use ncu to benchmark with this).\n"
                     "  -p,--path PATH        grayscale PGM image\n  -d,--debug_path DIR   existing directory for the stage dump\n"
                     "  -n,--iterations N     default 10\n";
        return 0;
    }
    const std::string img_path = cli::option(argc, argv, "-p", "--path");
    const std::string debug_path = cli::option(argc, argv, "-d", "--debug_path");
    const int iters = std::atoi(cli::option(argc, argv, "-n", "--iterations", "10").c_str());
    Imagef converted;
    if (!cli::read_pgm(img_path, converted)) {
        std::cout << "Image DNE" << std::endl;  // detection_example.cc:27-30
        return 1;
    }
    CudaSiftConfig config;
    config.upscale = false;
    config.col_width = int(converted.m_image_size.col);
    config.row_width = int(converted.m_image_size.row);
    std::cout << "r: " << config.row_width << ", c: " << config.col_width << std::endl;

    sift_cuda::Detector detector(config);
    if (!detector.gpuWarmUpAndAllocate()) return 1;
    if (!debug_path.empty()) detector.setDataGen(debug_path);
    cudaDeviceSynchronize();

    const auto t0 = std::chrono::steady_clock::now();
    for (int idx = 0; idx < iters; idx++) {
        detector.detectAndCompute(converted);
        cudaDeviceSynchronize();
    }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::cout << "keypoints: " << detector.total_size << ", " << ms / std::max(iters, 1) << " ms per detectAndCompute (host image, blocking)" << std::endl;
    return 0;
}
