// Test-infrastructure stub (oracle/_ref build only): shadows <GLFW/glfw3.h> for the reference's System.cpp.
// The viewer loop that uses these is never entered by the oracle wrapper (BENCHMARK / CPU only).
#pragma once
struct GLFWwindow;
inline int  glfwWindowShouldClose(GLFWwindow*) { return 1; }
inline void glfwPollEvents() {}
inline void glfwSwapBuffers(GLFWwindow*) {}
