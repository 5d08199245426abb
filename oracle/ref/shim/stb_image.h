// shim for stb_image (enginee/lib/init.sh:48-50): io.cpp's read_image is never reached with -DNO_RENDER.
#pragma once
inline unsigned char* stbi_load(const char*, int*, int*, int*, int) { return nullptr; }
inline void stbi_image_free(void*) {}
inline void stbi_set_flip_vertically_on_load(int) {}
