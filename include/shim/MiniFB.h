/* include/shim/MiniFB.h -- HEADLESS stand-in for the six MiniFB calls rayinteractive.c makes
 * (/root/reference/rayinteractive.c:118,122,181,183,191,199; API in dependencies/minifb/include/MiniFB.h:28-89).
 *
 * There is no window system on the GPU box, so the "window" is a frame counter: mfb_wait_sync() injects a
 * scripted list of key events through the registered keyboard callback (the same callback a real window
 * would call) and returns false after RTB_FRAMES frames; mfb_update_ex() time-stamps the presented frame.
 * At exit (mfb_timer_destroy) frame-time statistics are printed and the last frame can be written as a PNG.
 * Implemented in drivers/minifb_headless.c.  Key codes are MiniFB's own (MiniFB_enums.h:32-144).
 *
 *   RTB_FRAMES=240            number of frames before mfb_wait_sync returns false
 *   RTB_KEYS=RIGHT,D          events injected before every frame (UP DOWN LEFT RIGHT W A S D SPACE SHIFT)
 *   RTB_LAST_FRAME=out/x.png  write the last presented frame
 */
#ifndef RTB_SHIM_MINIFB_H
#define RTB_SHIM_MINIFB_H
#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    STATE_OK = 0,
    STATE_EXIT = -1,
    STATE_INVALID_WINDOW = -2,
    STATE_INVALID_BUFFER = -3,
    STATE_INTERNAL_ERROR = -4,
} mfb_update_state;

typedef enum {
    KB_KEY_UNKNOWN = -1,
    KB_KEY_SPACE = 32,
    KB_KEY_A = 65,
    KB_KEY_D = 68,
    KB_KEY_S = 83,
    KB_KEY_W = 87,
    KB_KEY_ESCAPE = 256,
    KB_KEY_RIGHT = 262,
    KB_KEY_LEFT = 263,
    KB_KEY_DOWN = 264,
    KB_KEY_UP = 265,
    KB_KEY_LEFT_SHIFT = 340,
} mfb_key;

typedef enum {
    KB_MOD_SHIFT = 0x0001,
    KB_MOD_CONTROL = 0x0002,
    KB_MOD_ALT = 0x0004,
    KB_MOD_SUPER = 0x0008,
    KB_MOD_CAPS_LOCK = 0x0010,
    KB_MOD_NUM_LOCK = 0x0020
} mfb_key_mod;

struct mfb_window;
struct mfb_timer;
typedef void (*mfb_keyboard_func)(struct mfb_window* window, mfb_key key, mfb_key_mod mod, bool isPressed);

struct mfb_window* mfb_open_ex(const char* title, unsigned width, unsigned height, unsigned flags);
void mfb_set_keyboard_callback(struct mfb_window* window, mfb_keyboard_func callback);
bool mfb_wait_sync(struct mfb_window* window);
mfb_update_state mfb_update_ex(struct mfb_window* window, void* buffer, unsigned width, unsigned height);
struct mfb_timer* mfb_timer_create(void);
void mfb_timer_destroy(struct mfb_timer* tmr);

#ifdef __cplusplus
}
#endif
#endif
