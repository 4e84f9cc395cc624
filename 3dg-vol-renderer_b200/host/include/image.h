// image.h -- float RGB image of the drop-in API (reference image.h:9-106): same constructors,
// accessors, 8-bit P6 reader/writer and RGBA export.  Extensions used by the GPU path: data()
// (the row-major 3*(y*W+x) float buffer vr_render writes straight into) and raw float dump/load.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <Eigen/Core>

class Image {
    unsigned int width = 0, height = 0;
    std::vector<float> pixels;  // 3 floats per pixel, row-major

    static unsigned char to_byte(float v) { return static_cast<unsigned char>(std::clamp(v * 255.0f, 0.0f, 255.0f)); }
    size_t at(unsigned int i, unsigned int j) const { return 3u * (size_t(j) * width + i); }

public:
    Image(unsigned int w, unsigned int h) : width(w), height(h), pixels(size_t(3) * w * h, 0.0f) {}

    // binary PPM, channels mapped to [0,1] by /255 (image.h:23-44)
    explicit Image(const std::string& filename) {
        std::ifstream in(filename, std::ios::binary);
        if (!in) throw std::runtime_error("Failed to open PPM file: " + filename);
        std::string magic;
        int maxval = 0;
        in >> magic;
        if (magic != "P6") throw std::runtime_error("Not a P6 PPM file.");
        in >> width >> height >> maxval;
        in.get();
        std::vector<unsigned char> raw(size_t(3) * width * height);
        in.read(reinterpret_cast<char*>(raw.data()), std::streamsize(raw.size()));
        pixels.resize(raw.size());
        for (size_t k = 0; k < raw.size(); ++k) pixels[k] = raw[k] / 255.f;
    }

    unsigned int get_width() const { return width; }
    unsigned int get_height() const { return height; }
    Eigen::Vector3f get_pixel(unsigned int i, unsigned int j) const {
        const size_t k = at(i, j);
        return Eigen::Vector3f(pixels[k], pixels[k + 1], pixels[k + 2]);
    }
    void set_pixel(unsigned int i, unsigned int j, const Eigen::Vector3f& rgb) {
        const size_t k = at(i, j);
        pixels[k] = rgb[0]; pixels[k + 1] = rgb[1]; pixels[k + 2] = rgb[2];
    }

    void make_PPM(const std::string& filename) const {
        std::vector<unsigned char> raw(pixels.size());
        for (size_t k = 0; k < pixels.size(); ++k) raw[k] = to_byte(pixels[k]);
        std::ofstream out(filename, std::ios::binary);
        out << "P6\n" << width << " " << height << "\n255\n";
        out.write(reinterpret_cast<const char*>(raw.data()), std::streamsize(raw.size()));
    }
    std::vector<uint8_t> get_rgba_buffer() const {
        std::vector<uint8_t> buf(size_t(4) * width * height);
        for (size_t p = 0; p < size_t(width) * height; ++p) {
            for (int c = 0; c < 3; ++c) buf[4 * p + c] = to_byte(pixels[3 * p + c]);
            buf[4 * p + 3] = 255;
        }
        return buf;
    }

    // ---- extensions
    float* data() { return pixels.data(); }
    const float* data() const { return pixels.data(); }
    void write_raw(const std::string& filename) const {  // 3*W*H float32, no header
        std::ofstream out(filename, std::ios::binary);
        out.write(reinterpret_cast<const char*>(pixels.data()), std::streamsize(pixels.size() * sizeof(float)));
    }
};
