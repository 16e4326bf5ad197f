/* Test infrastructure (oracle/_ref build only): minimal GL typedef stub.
 * The reference's types.h pulls <GLFW/glfw3.h> (glfw3.h:238 includes <GL/gl.h>) and use_cuda.h:15
 * pulls <cuda_gl_interop.h>; neither header is usable without these typedefs.  No GL function is
 * ever called by the oracle harness -- graphics APIs are unavailable on this box. */
#ifndef PTB200_ORACLE_GL_STUB_H
#define PTB200_ORACLE_GL_STUB_H
typedef unsigned int GLuint;
typedef int GLint;
typedef unsigned int GLenum;
typedef char GLchar;
typedef int GLsizei;
typedef float GLfloat;
typedef unsigned char GLboolean;
typedef unsigned int GLbitfield;
typedef void GLvoid;
typedef unsigned char GLubyte;
#define GL_TEXTURE_2D 0x0DE1
#endif
