// detectAndCompute on every image of a directory + matchBruteForce between consecutive images: the reference's
// tool/extract_and_match_example.cc:38-104 with the same option -d/--dir. Differences forced by this image:
// PGM files instead of cv::imread formats, and instead of cv::drawMatches + cv::imshow (needs a display) the
// matches go to stdout / --out DIR as text (matches_<i>.txt: "query train x0 y0 x1 y1") and a side-by-side PPM.
//   g++ -std=c++17 -I include -I tools/cli -I /usr/local/cuda/include tools/cli/extract_and_match_example.cc \
//       -L another_cuda_sift_b200 -lb200sift -L /usr/local/cuda/lib64 -lcudart -o extract_and_match_example
#include <algorithm>
#include <filesystem>
#include <fstream>
#include <iostream>

#include "image_io.hh"
#include "sift_cuda/interface/Detector.hh"
#include "sift_cuda/sift_func/Match.cuh"
#include "sift_cuda/types/CudaSiftConfig.hh"

namespace fs = std::filesystem;

int main(int argc, char** argv) {
    if (cli::flag(argc, argv, "-h", "--help")) {
        std::cout << "Runs detectAndCompute on each image of a directory and matches consecutive images.\n"
                     "  -d,--dir DIR   directory of equally sized grayscale .pgm images (sorted by name)\n"
                     "  -o,--out DIR   write matches_<i>.txt and matches_<i>.ppm there\n  -n,--num_features N   default 2000\n";
        return 0;
    }
    const fs::path img_dir(cli::option(argc, argv, "-d", "--dir"));
    const std::string out_dir = cli::option(argc, argv, "-o", "--out");
    if (!fs::exists(img_dir) || !fs::is_directory(img_dir)) {
        std::cout << "Directory does not exists." << std::endl;  // extract_and_match_example.cc:46-49
        return 1;
    }
    std::vector<std::string> paths;
    for (const auto& e : fs::directory_iterator(img_dir))
        if (e.path().extension() == ".pgm") paths.push_back(e.path().string());
    std::sort(paths.begin(), paths.end());
    std::vector<Imagef> host_img;
    for (const auto& p : paths) {
        Imagef im;
        if (cli::read_pgm(p, im)) host_img.push_back(im);
    }
    if (host_img.empty()) {
        std::cout << "No images." << std::endl;
        return 1;
    }
    CudaSiftConfig config;
    config.upscale = false;
    config.numFeatures = std::atoi(cli::option(argc, argv, "-n", "--num_features", "2000").c_str());
    config.col_width = int(host_img.front().m_image_size.col);
    config.row_width = int(host_img.front().m_image_size.row);
    sift_cuda::Detector detector(config);
    if (!detector.gpuWarmUpAndAllocate()) return 1;
    cudaDeviceSynchronize();

    thrust::host_vector<float3> prev_kpts;
    int prev_size = 0;
    for (size_t idx = 0; idx < host_img.size(); idx++) {
        detector.detectAndCompute(host_img[idx]);
        detector.copyToHost(false);
        cudaDeviceSynchronize();
        const int curr_size = detector.total_size;
        std::cout << paths[idx] << ": " << curr_size << " keypoints";
        if (idx != 0) {
            const std::vector<int> matches = sift_cuda::matchBruteForce(detector.prev_descriptor, prev_size, detector.device_descriptor, curr_size);
            const int good = int(std::count_if(matches.begin(), matches.end(), [](int m) { return m != -1; }));
            std::cout << ", " << good << " matches to the previous image";
            if (!out_dir.empty()) {
                std::ofstream txt(fs::path(out_dir) / ("matches_" + std::to_string(idx) + ".txt"));
                cli::Canvas canvas(host_img[idx - 1], host_img[idx]);
                const float xoff = host_img[idx - 1].m_image_size.col;
                for (int q = 0; q < int(matches.size()); q++) {
                    if (matches[q] == -1) continue;
                    const float3 a = prev_kpts[q], b = detector.final_kpts[matches[q]];
                    txt << q << " " << matches[q] << " " << a.x << " " << a.y << " " << b.x << " " << b.y << "\n";
                    canvas.line(a.x, a.y, b.x + xoff, b.y, (unsigned char)(60 + (q * 97) % 196), (unsigned char)(60 + (q * 57) % 196), 40);
                }
                canvas.write_ppm((fs::path(out_dir) / ("matches_" + std::to_string(idx) + ".ppm")).string());
            }
        }
        std::cout << std::endl;
        prev_kpts = detector.final_kpts;
        prev_size = curr_size;
    }
    return 0;
}
