// Mirror of the public surface of the reference's sift_cuda/types/HostImage.hh that the hot path touches:
// Size{row, col} (floats, HostImage.hh:21-28), Image<T> with m_data (shared_ptr<vector<T>>, row-major,
// unpadded) and m_image_size (HostImage.hh:36-45, 183-184), Imagef = Image<float> (HostImage.hh:188).
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <vector>

struct Size {
    float row;
    float col;
    Size() : row(0), col(0) {}
    Size(float r, float c) : row(r), col(c) {}
    Size(int r, int c) : row(float(r)), col(float(c)) {}
};

template <typename Data_T>
class Image {
  public:
    Image() {}
    Image(int rows, int cols) : m_data(std::make_shared<std::vector<Data_T>>(size_t(rows) * cols)), m_image_size(rows, cols) {}

    Data_T& at(int row, int col) {
        if (!m_data) throw std::runtime_error("at: Matrix accessed before initialization");  // HostImage.hh:43-45
        return m_data->at(size_t(row) * size_t(m_image_size.col) + col);
    }
    const Data_T& at(int row, int col) const {
        if (!m_data) throw std::runtime_error("at: Matrix accessed before initialization");
        return m_data->at(size_t(row) * size_t(m_image_size.col) + col);
    }

    std::shared_ptr<std::vector<Data_T>> m_data{};
    Size m_image_size{};
};

typedef Image<uint8_t> Image8U;
typedef Image<float> Imagef;
