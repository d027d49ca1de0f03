/* Link plumbing for the CPU oracle (oracle/build_ref.sh): no-op definitions of the OpenGL 1.x / GLU /
 * GLFW entry points the reference's interactive viewer code references.  None of them executes on the
 * headless "-f" path (reference src/main.cpp:175-188, src/application.cpp:48-50,184-187), which is the
 * only path the oracle drives.  This is not a graphics stand-in: every function is empty. */
#define STUB(name) void name(void) {}
STUB(glVertex3d) STUB(glEnable) STUB(glMatrixMode) STUB(glEnd) STUB(glBegin) STUB(glDisable)
STUB(glVertex3dv) STUB(glLineWidth) STUB(glColor4f) STUB(glVertex3i) STUB(glLoadIdentity)
STUB(glColor4fv) STUB(glVertex3f) STUB(glPushMatrix) STUB(glPopMatrix) STUB(glPointSize)
STUB(glVertex2f) STUB(glNormal3dv) STUB(glPushAttrib) STUB(glPopAttrib) STUB(glHint)
STUB(glViewport) STUB(glTranslatef) STUB(glTranslated) STUB(glScaled) STUB(glPolygonOffset)
STUB(glOrtho) STUB(glIsEnabled) STUB(glGetDoublev) STUB(glDepthMask) STUB(glColor3fv)
STUB(glGetIntegerv) STUB(glDrawPixels) STUB(gluPerspective) STUB(gluLookAt) STUB(glfwSetWindowSize)
