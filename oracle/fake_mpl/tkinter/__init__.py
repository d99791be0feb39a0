class Tk:
    def __init__(self, *a, **k):
        pass

    def withdraw(self):
        pass

    def destroy(self):
        pass

    def update(self):
        pass

    def attributes(self, *a, **k):
        pass
