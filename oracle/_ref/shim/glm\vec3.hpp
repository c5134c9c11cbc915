#include <glm/vec3.hpp>
