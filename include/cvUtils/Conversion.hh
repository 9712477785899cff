// OpenCV glue for the results of the SIFT hot path: same namespace and function names as the reference's
// cvUtils (cvUtils/Conversion.hh:12-71, Conversion.cc:10-59, ConversionImpl.hpp:7-82), header-only, so that
// code written against the reference (tool/extract_and_match_example.cc:60-97: cvMatToImage -> Detector ->
// localKptToCvKpt / cvtMatchToDMatch -> cv::drawMatches) keeps compiling with  -I <repo>/include.
//
// Needs OpenCV's core headers (<opencv2/core.hpp>: cv::Mat, cv::KeyPoint, cv::DMatch). This image has no
// OpenCV C++ development package; tests/shim/opencv2/core.hpp is a minimal stand-in used only to compile
// and exercise this header in tests/test_cvutils.py.
#pragma once
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <type_traits>
#include <vector>

#include <cuda_fp16.h>
#include <thrust/host_vector.h>

#include <opencv2/core.hpp>

#include "../sift_cuda/types/HostImage.hh"

namespace OpencvUtils {

// ConversionImpl.hpp:7-32: single-channel CV_8U / CV_32F / CV_64F Mat -> Image<T> (values are cast, not scaled:
// an 8-bit image becomes floats 0..255, which is what the detector's thresholds assume).
template <typename DataType_T>
Image<DataType_T> cvMatToImage(const cv::Mat& mat) {
    if (mat.channels() > 1) throw std::runtime_error("More channels than expected");
    Image<DataType_T> img{};
    img.m_image_size = Size{static_cast<float>(mat.rows), static_cast<float>(mat.cols)};
    std::vector<DataType_T> data;
    if (mat.type() == CV_64FC1) {
        data = std::vector<DataType_T>(mat.begin<double>(), mat.end<double>());
    } else if (mat.type() == CV_32FC1) {
        data = std::vector<DataType_T>(mat.begin<float>(), mat.end<float>());
    } else if (mat.type() == CV_8UC1) {
        data = std::vector<DataType_T>(mat.begin<uchar>(), mat.end<uchar>());
    }  // any other depth yields an empty image, like the reference
    img.m_data = std::make_shared<std::vector<DataType_T>>(std::move(data));
    return img;
}

// ConversionImpl.hpp:34-47: Image<uint8_t> / Image<float> -> CV_8UC1 / CV_32FC1 Mat (deep copy).
template <typename DataType_T>
cv::Mat imageToCvMat(Image<DataType_T> image) {
    static_assert(std::is_same_v<DataType_T, uint8_t> || std::is_same_v<DataType_T, float>, "8-bit or float images only");
    const int cv_data_type = std::is_same_v<DataType_T, uint8_t> ? CV_8UC1 : CV_32FC1;
    cv::Mat newMat(int(image.m_image_size.row), int(image.m_image_size.col), cv_data_type);
    std::memcpy(newMat.data, image.m_data->data(), image.m_data->size() * sizeof(DataType_T));
    return newMat;
}

// Conversion.cc:10-20: same shape, channel count and bits.
inline bool areEqual(const cv::Mat& a, const cv::Mat& b) {
    if (a.channels() != b.channels() || a.rows != b.rows || a.cols != b.cols || a.type() != b.type()) return false;
    for (int r = 0; r < a.rows; r++)
        if (std::memcmp(a.ptr(r), b.ptr(r), size_t(a.cols) * a.elemSize()) != 0) return false;
    return true;
}

// ConversionImpl.hpp:50-64: min-max stretch to 0..255, truncating cast.
template <typename DataType_T>
Image8U normalize(const Image<DataType_T>& image) {
    const DataType_T lo = *std::min_element(image.m_data->begin(), image.m_data->end());
    const DataType_T hi = *std::max_element(image.m_data->begin(), image.m_data->end());
    Image8U out{};
    out.m_image_size = image.m_image_size;
    out.m_data = std::make_shared<std::vector<uint8_t>>();
    out.m_data->reserve(image.m_data->size());
    const DataType_T scale = 255 / (hi - lo);
    for (const DataType_T v : *image.m_data) out.m_data->push_back(static_cast<uint8_t>((v - lo) * scale));
    return out;
}

// Conversion.cc:22-45: pt = (x, y) in full-resolution pixels, octave = the packed octave word (float -> int),
// response = |contrast|, size, angle in degrees. size = -1 converts every entry of `kpts`; callers pass
// Detector::total_size (the host mirrors are numFeatures long).
inline std::vector<cv::KeyPoint> localKptToCvKpt(const thrust::host_vector<float3>& kpts,
                                                 const thrust::host_vector<float4>& features, int size = -1) {
    assert(kpts.size() == features.size());
    const size_t n = size == -1 ? kpts.size() : size_t(size);
    std::vector<cv::KeyPoint> out;
    out.reserve(n);
    for (size_t i = 0; i < n; i++) {
        cv::KeyPoint k;
        k.pt.x = kpts[i].x;
        k.pt.y = kpts[i].y;
        k.octave = int(features[i].x);
        k.response = features[i].z;
        k.size = features[i].y;
        k.angle = features[i].w;
        out.push_back(k);
    }
    return out;
}

// ConversionImpl.hpp:66-82: row i = descriptor of keypoint i as 128 floats (half values 0..512 widened).
template <typename Data_T>
cv::Mat descriptorToCvMat(const thrust::host_vector<Data_T>& descriptors, int num_pts) {
    num_pts = std::min(num_pts, int(descriptors.size() / 128));
    cv::Mat desc(num_pts, 128, CV_32FC1);
    for (int r = 0; r < num_pts; r++) {
        float* row = desc.template ptr<float>(r);
        for (int c = 0; c < 128; c++) {
            if constexpr (std::is_same_v<Data_T, half>) {
                row[c] = __half2float(descriptors[size_t(r) * 128 + c]);
            } else {
                row[c] = static_cast<float>(descriptors[size_t(r) * 128 + c]);
            }
        }
    }
    return desc;
}

// Conversion.cc:47-58: query index -> train index for every accepted match (-1 = none); distance is not returned.
inline std::vector<cv::DMatch> cvtMatchToDMatch(const std::vector<int>& match) {
    std::vector<cv::DMatch> out;
    out.reserve(match.size());
    for (int i = 0; i < int(match.size()); i++)
        if (match[i] != -1) out.push_back(cv::DMatch(i, match[i], 0.f));
    return out;
}

}  // namespace OpencvUtils
