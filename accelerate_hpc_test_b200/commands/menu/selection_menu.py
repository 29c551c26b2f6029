"""Arrow-key bullet menu for the questionnaire (reference commands/menu/*, ~500 LoC over five files).

One file here: raw-mode key reading, ANSI cursor movement and the menu loop.  Falls back to a
numbered prompt when stdin is not a TTY (pipes, CI, notebooks).
"""

import os
import sys

CURSOR_UP, CURSOR_DOWN = "\033[{}A", "\033[{}B"
CLEAR_LINE = "\033[2K\r"
KEY_UP, KEY_DOWN, KEY_ENTER, KEY_INTERRUPT = "up", "down", "enter", "interrupt"


def _read_key():
    """One key press from a raw TTY, decoded to 'up' / 'down' / 'enter' / 'interrupt' / a character."""
    if os.name == "nt":
        import msvcrt

        ch = msvcrt.getch()
        if ch in (b"\x00", b"\xe0"):
            ch2 = msvcrt.getch()
            return {b"H": KEY_UP, b"P": KEY_DOWN}.get(ch2, "")
        ch = ch.decode(errors="ignore")
    else:
        import termios
        import tty

        fd = sys.stdin.fileno()
        old = termios.tcgetattr(fd)
        try:
            tty.setraw(fd)
            ch = sys.stdin.read(1)
            if ch == "\x1b":
                seq = sys.stdin.read(2)
                return {"[A": KEY_UP, "[B": KEY_DOWN}.get(seq, "")
        finally:
            termios.tcsetattr(fd, termios.TCSADRAIN, old)
    if ch in ("\r", "\n"):
        return KEY_ENTER
    if ch == "\x03":
        return KEY_INTERRUPT
    return ch


class BulletMenu:
    """``BulletMenu(prompt, choices).run(default_choice=0)`` returns the index of the chosen entry."""

    def __init__(self, prompt: str = None, choices: list = []):
        self.position = 0
        self.choices = choices
        self.prompt = prompt
        self.arrow_char = "➔ " if sys.platform != "win32" else "*"

    def _draw(self, first: bool):
        out = sys.stdout
        if not first:
            out.write(CURSOR_UP.format(len(self.choices)))
        for i, choice in enumerate(self.choices):
            out.write(CLEAR_LINE)
            if i == self.position:
                out.write(f" {self.arrow_char}\033[32m{choice}\033[0m\n")
            else:
                out.write(f"    {choice}\n")
        out.flush()

    def _fallback(self, default_choice):
        print(self.prompt)
        for i, c in enumerate(self.choices):
            print(f"  [{i}] {c}")
        while True:
            raw = input(f"Please select a choice using the number [{default_choice}]: ").strip()
            if raw == "":
                return default_choice
            try:
                v = int(raw)
                if 0 <= v < len(self.choices):
                    return v
            except ValueError:
                pass
            print("Please enter a valid number.")

    def run(self, default_choice: int = 0):
        if not sys.stdin.isatty() or os.environ.get("ACCELERATE_DISABLE_RICH_MENU", "") == "1":
            return self._fallback(default_choice)
        if self.prompt:
            print(self.prompt)
            print("Please select a choice using the arrow or number keys, and selecting with enter")
        self.position = default_choice
        self._draw(first=True)
        try:
            while True:
                key = _read_key()
                if key == KEY_UP:
                    self.position = (self.position - 1) % len(self.choices)
                elif key == KEY_DOWN:
                    self.position = (self.position + 1) % len(self.choices)
                elif key == KEY_ENTER:
                    return self.position
                elif key == KEY_INTERRUPT:
                    raise KeyboardInterrupt
                elif isinstance(key, str) and key.isdigit() and int(key) < len(self.choices):
                    self.position = int(key)
                self._draw(first=False)
        finally:
            sys.stdout.flush()
