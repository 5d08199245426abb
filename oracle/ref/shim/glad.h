// shim for the generated glad loader (enginee/lib/init.sh:28-29): types only.  The hot-path TUs are built with
// -DNO_RENDER and never call OpenGL; headers they include still mention GL types.  TEST INFRASTRUCTURE ONLY.
#pragma once
typedef unsigned int GLuint;
typedef int GLint;
typedef unsigned int GLenum;
typedef int GLsizei;
typedef float GLfloat;
typedef unsigned char GLboolean;
typedef char GLchar;
typedef unsigned char GLubyte;
typedef unsigned int GLbitfield;
typedef double GLdouble;
