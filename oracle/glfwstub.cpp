// Test infrastructure (oracle/_ref build only): link stubs for the three GLFW entry points the
// reference's host headers reference (bvhBuilder.h:37,264 / scene.h:339 glfwGetTime; types.h:612-637
// and keyboard.h glfwGetKey; keyboard.h glfwSetInputMode).  No window system exists on this box.
#include <chrono>
extern "C" {
double glfwGetTime(void) {
    using namespace std::chrono;
    static const auto t0 = steady_clock::now();
    return duration<double>(steady_clock::now() - t0).count();
}
int glfwGetKey(void*, int) { return 0; }
void glfwSetInputMode(void*, int, int) {}
}
