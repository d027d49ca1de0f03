/* Build plumbing for the CPU oracle only: the reference's vendored glew.h includes <GL/glu.h>,
 * which this image does not have.  Nothing declared here is ever called on the headless (-f) path
 * (reference src/main.cpp:175-188). */
#ifndef RRT_ORACLE_GLU_STUB_H
#define RRT_ORACLE_GLU_STUB_H
#ifdef __cplusplus
extern "C" {
#endif
void gluLookAt(double, double, double, double, double, double, double, double, double);
void gluPerspective(double, double, double, double);
#ifdef __cplusplus
}
#endif
#endif
