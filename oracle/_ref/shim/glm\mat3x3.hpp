#include <glm/mat3x3.hpp>
