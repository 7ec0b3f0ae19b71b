// image_out.hpp — what `RgbImage::save("output.png")` does for the reference (src/main.rs:66): 8-bit RGB PNG (zlib
// deflate, filter 0) or binary PPM, chosen by the file extension.  rgb8: width*height*3 bytes, row 0 = top.
#pragma once
#include <cstdint>
#include <string>

namespace rtbhost {

// Throws std::runtime_error on I/O failure.
void save_image(const std::string& path, const uint8_t* rgb8, int width, int height);
void write_png(const std::string& path, const uint8_t* rgb8, int width, int height);
void write_ppm(const std::string& path, const uint8_t* rgb8, int width, int height);

}  // namespace rtbhost
