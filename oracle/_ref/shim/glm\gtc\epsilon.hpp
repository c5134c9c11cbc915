#include <glm/gtc/epsilon.hpp>
