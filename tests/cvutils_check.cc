// Exercises include/cvUtils/Conversion.hh against tests/shim/opencv2/core.hpp (host-side only, no GPU).
#include <cstdio>
#include "cvUtils/Conversion.hh"

int main() {
    int bad = 0;
    cv::Mat m(3, 4, CV_8UC1);
    for (int i = 0; i < 12; i++) m.data[i] = uchar(i * 20);
    Imagef img = OpencvUtils::cvMatToImage<float>(m);
    bad += !(img.m_image_size.row == 3.f && img.m_image_size.col == 4.f && img.m_data->size() == 12 && img.at(2, 3) == 220.f);
    cv::Mat back = OpencvUtils::imageToCvMat<float>(img);
    bad += !(back.type() == CV_32FC1 && back.at<float>(1, 2) == 120.f);
    Image8U n8 = OpencvUtils::normalize<float>(img);
    bad += !(n8.m_data->front() == 0 && n8.m_data->back() == 255);
    cv::Mat m8 = OpencvUtils::imageToCvMat<uint8_t>(n8);
    bad += !OpencvUtils::areEqual(m8, m8) || OpencvUtils::areEqual(m8, back);

    thrust::host_vector<float3> k(3);
    thrust::host_vector<float4> f(3);
    k[1] = make_float3(10.f, 20.f, 2.f);
    f[1] = make_float4(float(1 + (2 << 8) + (255 << 16)), 3.5f, 0.05f, 270.f);
    auto cvk = OpencvUtils::localKptToCvKpt(k, f, 2);
    bad += !(cvk.size() == 2 && cvk[1].pt.x == 10.f && cvk[1].pt.y == 20.f && cvk[1].octave == 1 + (2 << 8) + (255 << 16) &&
             cvk[1].size == 3.5f && cvk[1].response == 0.05f && cvk[1].angle == 270.f);
    bad += !(OpencvUtils::localKptToCvKpt(k, f).size() == 3);

    thrust::host_vector<half> d(2 * 128);
    for (int i = 0; i < 256; i++) d[i] = __float2half(float(i));
    cv::Mat dm = OpencvUtils::descriptorToCvMat<half>(d, 5);  // clamped to the 2 rows present
    bad += !(dm.rows == 2 && dm.cols == 128 && dm.at<float>(1, 127) == 255.f && dm.at<float>(0, 1) == 1.f);

    auto dmatch = OpencvUtils::cvtMatchToDMatch({4, -1, 0, -1, 2});
    bad += !(dmatch.size() == 3 && dmatch[1].queryIdx == 2 && dmatch[1].trainIdx == 0 && dmatch[2].queryIdx == 4 && dmatch[2].trainIdx == 2);
    std::printf("cvutils_check: %s\n", bad ? "FAIL" : "ok");
    return bad;
}
