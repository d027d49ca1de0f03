// Link plumbing for the CPU oracle (oracle/build_ref.sh): empty bodies for the two CGL classes whose
// real sources need Freetype/GLFW (CGL/src/osdtext.cpp, CGL/src/viewer.cpp are not compiled).
// Application holds an OSDText member, so its constructor/destructor must link; nothing else is reached
// on the headless path.
#include "CGL/osdtext.h"
#include "CGL/viewer.h"
namespace CGL {
OSDText::OSDText() {}
OSDText::~OSDText() {}
int OSDText::init(bool) { return 0; }
void OSDText::render() {}
void OSDText::clear() {}
void OSDText::resize(size_t, size_t) {}
int OSDText::add_line(float, float, std::string, size_t, Color) { return 0; }
Viewer::Viewer(void) {}
Viewer::~Viewer(void) {}
void Viewer::init(void) {}
void Viewer::start(void) {}
void Viewer::set_renderer(Renderer*) {}
void Viewer::resize_callback(GLFWwindow*, int, int) {}
GLFWwindow* Viewer::window = nullptr;
}
