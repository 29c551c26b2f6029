from .selection_menu import BulletMenu
