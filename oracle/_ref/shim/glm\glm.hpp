#include <glm/glm.hpp>
