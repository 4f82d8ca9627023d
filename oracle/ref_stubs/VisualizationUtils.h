// Test-infrastructure stub (oracle/_ref build only): shadows simulator/visualization/VisualizationUtils.h.
// Only the names System.cpp mentions are declared; every function is an inert no-op because the
// viewer ("unavailable on this box") is never entered by the oracle wrapper.
#pragma once
#include <vector_types.h>
struct GLFWwindow;
class ParticleRenderer {
public:
    enum class ColorMode { VELOCITY_MAGNITUDE };
    void  setColorMode(ColorMode) {}
    void  resetView() {}
    void  setPointSize(float) {}
    void  updatePositions(const double4*, const double4*, int, float, float, float, float) {}
    float getTranslateX() const { return 0.f; }
    float getTranslateY() const { return 0.f; }
    float getZoom() const { return 1.f; }
    void  render() {}
};
namespace VisualizationUtils {
    inline float scale = 1.0f;
    inline ParticleRenderer* renderer = nullptr;
    inline GLFWwindow* window = nullptr;
    inline void printKeyboardControls() {}
    inline bool initVisualization(int, const char*, bool) { return false; }
    inline void setupKeyboardCallback() {}
    inline void updateWindowTitle(int, int, double, double) {}
    inline void cleanupVisualization() {}
}
